// comm.cu -- multi-GPU plumbing: one process per GPU, NCCL over NVLink 5 / NVSwitch.
//
// The path shards by slabs of the slowest grid index: a rank owns planes
// [j0, j1) of every species.  Only two exchange patterns exist:
//   * halo: one plane per species to each slab neighbour (ncclSend/ncclRecv in one group) --
//     before the batched residual (y) and before every SpMV (the Krylov direction);
//   * scalars: ncclAllReduce of 1-2 doubles per fused dot / norm, and of the Newton test values.
// Messages are small (a 256^2 plane is 512 KiB), so the cost is latency, not bandwidth: all sends
// and receives of one exchange go into a single NCCL group, on the solver's stream, with no host
// synchronisation.
#include <cstdlib>
#include <vector>

#include "solver.cuh"

#ifdef RS_WITH_NCCL
#include <nccl.h>
#endif

namespace rs {

#ifdef RS_WITH_NCCL
static ncclComm_t g_comm = nullptr;
static int g_nranks = 1, g_rank = 0;

#define RS_NCCL(call)                                                                       \
    do {                                                                                    \
        ncclResult_t r_ = (call);                                                           \
        if (r_ != ncclSuccess) {                                                            \
            set_error("%s failed: %s", #call, ncclGetErrorString(r_));                      \
            return RS_ECOMM;                                                                \
        }                                                                                   \
    } while (0)

bool comm_active() { return g_comm != nullptr && g_nranks > 1; }
int comm_nranks() { return g_nranks; }
int comm_rank() { return g_rank; }

// ---- NVLink peer mailboxes (see PeerRed in rs_internal.cuh) -------------------------------------
static double *g_mail_local = nullptr;
static double *g_mail[PEER_MAX] = {nullptr};
static bool g_peer_ok = false;
static unsigned *g_seq_dev = nullptr;   // [0] all-reduces, [1] halo pushes completed on this rank (device counters)

bool comm_peer_enabled() { return g_peer_ok && comm_active(); }

PeerRed comm_next_peer_red()
{
    PeerRed pr;
    if (!comm_peer_enabled()) return pr;
    pr.nranks = g_nranks;
    pr.rank = g_rank;
    pr.seq = g_seq_dev;
    for (int p = 0; p < g_nranks; p++) pr.mail[p] = g_mail[p];
    return pr;
}

// Maps every rank's mailbox into every process (CUDA IPC over NVLink P2P).  Any failure simply leaves
// the NCCL all-reduce path in charge.
static void peer_setup()
{
    g_peer_ok = false;
    if (g_nranks > PEER_MAX || getenv("RS_NO_PEER")) return;
    const size_t bytes = sizeof(double) * (2 * PEER_MAX * PEER_SLOT + 8);   // reduction slots + halo flag words
    if (cudaMalloc((void **)&g_mail_local, bytes) != cudaSuccess) { cudaGetLastError(); return; }
    cudaMemset(g_mail_local, 0, bytes);
    cudaIpcMemHandle_t mine;
    if (cudaIpcGetMemHandle(&mine, g_mail_local) != cudaSuccess) { cudaGetLastError(); return; }
    char *d_h = nullptr;
    const size_t hs = sizeof(cudaIpcMemHandle_t);
    if (cudaMalloc((void **)&d_h, hs * g_nranks) != cudaSuccess) { cudaGetLastError(); return; }
    cudaMemcpy(d_h + hs * g_rank, &mine, hs, cudaMemcpyHostToDevice);
    std::vector<cudaIpcMemHandle_t> all(g_nranks);
    int ok = 1;
    if (ncclAllGather(d_h + hs * g_rank, d_h, hs, ncclChar, g_comm, nullptr) != ncclSuccess) ok = 0;
    if (ok && cudaStreamSynchronize(nullptr) != cudaSuccess) ok = 0;
    if (ok) cudaMemcpy(all.data(), d_h, hs * g_nranks, cudaMemcpyDeviceToHost);
    for (int p = 0; ok && p < g_nranks; p++) {
        if (p == g_rank) { g_mail[p] = g_mail_local; continue; }
        void *ptr = nullptr;
        if (cudaIpcOpenMemHandle(&ptr, all[p], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
        g_mail[p] = (double *)ptr;
    }
    // everybody must agree, and nobody may use a mailbox before all of them are mapped
    int *d_ok = (int *)d_h;
    cudaMemcpy(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice);
    if (ncclAllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, g_comm, nullptr) == ncclSuccess && cudaStreamSynchronize(nullptr) == cudaSuccess)
        cudaMemcpy(&ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost);
    else
        ok = 0;
    cudaFree(d_h);
    g_peer_ok = ok != 0;
    if (ok) { ok = cudaMalloc((void **)&g_seq_dev, 64) == cudaSuccess; if (ok) cudaMemset(g_seq_dev, 0, 64); else cudaGetLastError(); g_peer_ok = ok != 0; }
}

// ---- halo push: neighbours' vectors mapped for direct stores --------------------------------------


// Maps the slab neighbours' copies of `count` vectors (allocation bases) into this process.
// lo[i] / hi[i] receive the lower / upper neighbour's pointer for vector i (nullptr at the ends).
// Collective over all ranks.  Returns false (and leaves the NCCL halo exchange in charge) on any failure.
bool comm_share_vectors(void *const *mine, int count, void **lo, void **hi)
{
    for (int i = 0; i < count; i++) { lo[i] = nullptr; hi[i] = nullptr; }
    if (!comm_peer_enabled() || count > 4) return false;
    const size_t hs = sizeof(cudaIpcMemHandle_t);
    std::vector<cudaIpcMemHandle_t> own(count), all((size_t)count * g_nranks);
    int ok = 1;
    for (int i = 0; i < count; i++)
        if (cudaIpcGetMemHandle(&own[i], mine[i]) != cudaSuccess) { cudaGetLastError(); ok = 0; }
    char *d_h = nullptr;
    if (cudaMalloc((void **)&d_h, hs * count * g_nranks + 64) != cudaSuccess) { cudaGetLastError(); return false; }
    cudaMemcpy(d_h + hs * count * g_rank, own.data(), hs * count, cudaMemcpyHostToDevice);
    if (ncclAllGather(d_h + hs * count * g_rank, d_h, hs * count, ncclChar, g_comm, nullptr) != ncclSuccess) ok = 0;
    if (cudaStreamSynchronize(nullptr) != cudaSuccess) ok = 0;
    if (ok) cudaMemcpy(all.data(), d_h, hs * count * g_nranks, cudaMemcpyDeviceToHost);
    for (int i = 0; ok && i < count; i++) {
        if (g_rank > 0 && cudaIpcOpenMemHandle(&lo[i], all[(size_t)(g_rank - 1) * count + i], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); lo[i] = nullptr; ok = 0; }
        if (g_rank < g_nranks - 1 && cudaIpcOpenMemHandle(&hi[i], all[(size_t)(g_rank + 1) * count + i], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); hi[i] = nullptr; ok = 0; }
    }
    int *d_ok = (int *)d_h;
    cudaMemcpy(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice);
    if (ncclAllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, g_comm, nullptr) == ncclSuccess && cudaStreamSynchronize(nullptr) == cudaSuccess)
        cudaMemcpy(&ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost);
    else
        ok = 0;
    cudaFree(d_h);
    if (!ok) comm_unshare_vectors(count, lo, hi);
    return ok != 0;
}

void comm_unshare_vectors(int count, void **lo, void **hi)
{
    for (int i = 0; i < count; i++) {
        if (lo[i]) cudaIpcCloseMemHandle(lo[i]);
        if (hi[i]) cudaIpcCloseMemHandle(hi[i]);
        lo[i] = hi[i] = nullptr;
    }
}

// Pool of two peer-mapped vectors, grown on demand, released with the communicator.
static size_t g_pool_n = 0;
static double *g_pool_mine[2] = {nullptr, nullptr};
static void *g_pool_lo[2] = {nullptr, nullptr}, *g_pool_hi[2] = {nullptr, nullptr};
static bool g_pool_ok = false;

// bumped whenever the pool or the arena below is re-created: a solver handle that still holds pointers into the old
// buffers must not run on them (krylov_solve checks)
static int g_pool_generation = 0;
int comm_pool_generation() { return g_pool_generation; }

static void pool_release()
{
    g_pool_generation++;
    if (g_pool_ok) comm_unshare_vectors(2, g_pool_lo, g_pool_hi);
    for (int i = 0; i < 2; i++) { if (g_pool_mine[i]) cudaFree(g_pool_mine[i]); g_pool_mine[i] = nullptr; }
    g_pool_n = 0;
    g_pool_ok = false;
}

bool comm_pooled_vectors(size_t n, double **mine, void **lo, void **hi)
{
    if (!comm_peer_enabled()) return false;
    if (!g_pool_ok || g_pool_n < n) {   // every rank asks for the same n in the same order: the branch is collective
        cudaDeviceSynchronize();
        pool_release();
        int ok = 1;
        for (int i = 0; i < 2; i++)
            if (cudaMalloc((void **)&g_pool_mine[i], (n + 8) * sizeof(double)) != cudaSuccess) { cudaGetLastError(); g_pool_mine[i] = nullptr; ok = 0; }
        for (int i = 0; ok && i < 2; i++) cudaMemset(g_pool_mine[i], 0, (n + 8) * sizeof(double));
        void *m[2] = {g_pool_mine[0], g_pool_mine[1]};
        // the share is collective even if this rank failed to allocate: pass what we have, it reports failure for all
        g_pool_ok = (g_pool_mine[0] && g_pool_mine[1]) ? comm_share_vectors(m, 2, g_pool_lo, g_pool_hi) : false;
        if (!g_pool_ok) { pool_release(); return false; }
        g_pool_n = n;
    }
    for (int i = 0; i < 2; i++) { mine[i] = g_pool_mine[i]; lo[i] = g_pool_lo[i]; hi[i] = g_pool_hi[i]; }
    return true;
}

// One more peer-mapped buffer, for the level vectors of the multigrid hierarchy whose halo planes travel (mg.cu):
// same life cycle as the pool above.
static size_t g_arena_n = 0;
static double *g_arena_mine = nullptr;
static void *g_arena_lo = nullptr, *g_arena_hi = nullptr;
static bool g_arena_ok = false;

static void arena_release()
{
    g_pool_generation++;
    if (g_arena_ok) comm_unshare_vectors(1, &g_arena_lo, &g_arena_hi);
    if (g_arena_mine) cudaFree(g_arena_mine);
    g_arena_mine = nullptr; g_arena_lo = g_arena_hi = nullptr;
    g_arena_n = 0;
    g_arena_ok = false;
}

bool comm_pooled_arena(size_t n, double **mine, void **lo, void **hi)
{
    if (!comm_peer_enabled()) return false;
    if (!g_arena_ok || g_arena_n < n) {   // collective: every rank asks for the same n in the same order
        cudaDeviceSynchronize();
        arena_release();
        if (cudaMalloc((void **)&g_arena_mine, (n + 8) * sizeof(double)) != cudaSuccess) { cudaGetLastError(); g_arena_mine = nullptr; }
        if (g_arena_mine) cudaMemset(g_arena_mine, 0, (n + 8) * sizeof(double));
        void *m[1] = {g_arena_mine};
        // every rank takes part in the share even if its own allocation failed: a NULL handle fails it for all
        g_arena_ok = comm_share_vectors(m, 1, &g_arena_lo, &g_arena_hi);
        if (!g_arena_ok) { arena_release(); return false; }
        g_arena_n = n;
    }
    *mine = g_arena_mine; *lo = g_arena_lo; *hi = g_arena_hi;
    return true;
}

// flag words of the halo handshake live behind the reduction slots of every mailbox:
//   [HALO_FROM_LO] is written by rank-1, [HALO_FROM_HI] by rank+1
HaloSignal comm_next_halo_signal()
{
    HaloSignal hs;
    if (!comm_peer_enabled()) return hs;
    hs.seq = g_seq_dev + 1;
    const size_t off = (size_t)2 * PEER_MAX * PEER_SLOT;
    hs.wait_lo = g_rank > 0 ? g_mail_local + off + 0 : nullptr;
    hs.wait_hi = g_rank < g_nranks - 1 ? g_mail_local + off + 1 : nullptr;
    hs.tell_lo = g_rank > 0 ? g_mail[g_rank - 1] + off + 1 : nullptr;              // I am their upper neighbour
    hs.tell_hi = g_rank < g_nranks - 1 ? g_mail[g_rank + 1] + off + 0 : nullptr;   // I am their lower neighbour
    return hs;
}

int comm_allreduce_sum(Ctx &c, double *d_buf, int count)
{
    if (!comm_active()) return RS_OK;
    RS_NCCL(ncclAllReduce(d_buf, d_buf, count, ncclDouble, ncclSum, g_comm, c.stream));
    return RS_OK;
}

int comm_allreduce_max(Ctx &c, double *d_buf, int count)
{
    if (!comm_active()) return RS_OK;
    RS_NCCL(ncclAllReduce(d_buf, d_buf, count, ncclDouble, ncclMax, g_comm, c.stream));
    return RS_OK;
}

int comm_halo_exchange(Ctx &c, double *v, const HaloPlan &h)
{
    if (!comm_active() || h.nspec == 0) return RS_OK;
    c.launches++;
    c.count[K_HALO]++;
    RS_NCCL(ncclGroupStart());
    for (int s = 0; s < h.nspec; s++) {
        const long long base = (long long)s * h.ncell;
        if (h.has_lo) {
            RS_NCCL(ncclSend(v + base + (long long)h.j0 * h.plane, h.plane, ncclDouble, g_rank - 1, g_comm, c.stream));
            RS_NCCL(ncclRecv(v + base + (long long)(h.j0 - 1) * h.plane, h.plane, ncclDouble, g_rank - 1, g_comm, c.stream));
        }
        if (h.has_hi) {
            RS_NCCL(ncclSend(v + base + (long long)(h.j1 - 1) * h.plane, h.plane, ncclDouble, g_rank + 1, g_comm, c.stream));
            RS_NCCL(ncclRecv(v + base + (long long)h.j1 * h.plane, h.plane, ncclDouble, g_rank + 1, g_comm, c.stream));
        }
    }
    RS_NCCL(ncclGroupEnd());
    return RS_OK;
}

// every rank ends up with the whole vector: each slab is broadcast from its owner
int comm_gather_slabs(Ctx &c, double *v, int nspec, int ncell, int plane, int d1)
{
    if (!comm_active()) return RS_OK;
    RS_NCCL(ncclGroupStart());
    for (int r = 0; r < g_nranks; r++) {
        int j0, j1;
        rs_slab_bounds(d1, g_nranks, r, &j0, &j1);
        for (int s = 0; s < nspec; s++) {
            double *p = v + (long long)s * ncell + (long long)j0 * plane;
            RS_NCCL(ncclBroadcast(p, p, (size_t)(j1 - j0) * plane, ncclDouble, r, g_comm, c.stream));
        }
    }
    RS_NCCL(ncclGroupEnd());
    return RS_OK;
}
#else
bool comm_active() { return false; }
bool comm_peer_enabled() { return false; }
PeerRed comm_next_peer_red() { return PeerRed(); }
bool comm_share_vectors(void *const *, int count, void **lo, void **hi) { for (int i = 0; i < count; i++) lo[i] = hi[i] = nullptr; return false; }
bool comm_pooled_vectors(size_t, double **, void **, void **) { return false; }
bool comm_pooled_arena(size_t, double **, void **, void **) { return false; }
int comm_pool_generation() { return 0; }
void comm_unshare_vectors(int, void **, void **) {}
HaloSignal comm_next_halo_signal() { return HaloSignal(); }
int comm_nranks() { return 1; }
int comm_rank() { return 0; }
int comm_allreduce_sum(Ctx &, double *, int) { return RS_OK; }
int comm_allreduce_max(Ctx &, double *, int) { return RS_OK; }
int comm_halo_exchange(Ctx &, double *, const HaloPlan &) { return RS_OK; }
int comm_gather_slabs(Ctx &, double *, int, int, int, int) { return RS_OK; }
#endif

}  // namespace rs

extern "C" int rs_comm_unique_id(char id[128])
{
#ifdef RS_WITH_NCCL
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is expected to be 128 bytes");
    ncclUniqueId u;
    ncclResult_t r = ncclGetUniqueId(&u);
    if (r != ncclSuccess) { rs::set_error("ncclGetUniqueId failed: %s", ncclGetErrorString(r)); return RS_ECOMM; }
    memcpy(id, &u, 128);
    return RS_OK;
#else
    (void)id;
    rs::set_error("built without NCCL");
    return RS_ECOMM;
#endif
}

extern "C" int rs_comm_init(int nranks, int rank, const char id[128])
{
#ifdef RS_WITH_NCCL
    using namespace rs;
    if (g_comm) { set_error("communicator already initialised"); return RS_ECOMM; }
    if (nranks < 1 || rank < 0 || rank >= nranks) { set_error("bad rank/nranks"); return RS_EINVAL; }
    RS_TRY(ensure_device());
    ncclUniqueId u;
    memcpy(&u, id, 128);
    ncclResult_t r = ncclCommInitRank(&g_comm, nranks, u, rank);
    if (r != ncclSuccess) { g_comm = nullptr; set_error("ncclCommInitRank failed: %s", ncclGetErrorString(r)); return RS_ECOMM; }
    g_nranks = nranks;
    g_rank = rank;
    peer_setup();
    return RS_OK;
#else
    (void)nranks; (void)rank; (void)id;
    rs::set_error("built without NCCL");
    return RS_ECOMM;
#endif
}

extern "C" int rs_comm_nranks(void) { return rs::comm_nranks(); }
extern "C" int rs_comm_rank(void) { return rs::comm_rank(); }

extern "C" void rs_comm_finalize(void)
{
#ifdef RS_WITH_NCCL
    if (rs::g_comm) {
        cudaDeviceSynchronize();
        for (int p = 0; p < rs::PEER_MAX; p++) {
            if (rs::g_mail[p] && rs::g_mail[p] != rs::g_mail_local) cudaIpcCloseMemHandle(rs::g_mail[p]);
            rs::g_mail[p] = nullptr;
        }
        rs::pool_release();
        rs::arena_release();
        if (rs::g_seq_dev) { cudaFree(rs::g_seq_dev); rs::g_seq_dev = nullptr; }
        if (rs::g_mail_local) { cudaFree(rs::g_mail_local); rs::g_mail_local = nullptr; }
        rs::g_peer_ok = false;
        ncclCommDestroy(rs::g_comm);
        rs::g_comm = nullptr; rs::g_nranks = 1; rs::g_rank = 0;
    }
#endif
}

// Planes [j0, j1) of the slowest grid dimension owned by `rank`: as even as possible (the first d1 % nranks ranks get
// one more), with the interior bounds then moved down to a multiple of A = 2, 4, 8 or 16 planes while every rank keeps
// at least 3A: whole 2^k aggregates per rank are what lets the finest multigrid levels span the ranks (mg.cu).
static int slab_bound(int d1, int nranks, int r)
{
    if (r <= 0) return 0;
    if (r >= nranks) return d1;
    const int base = d1 / nranks, rem = d1 % nranks;
    int A = 1;
    while (A < 16 && d1 >= nranks * 4 * (2 * A)) A *= 2;
    const int even = r * base + (r < rem ? r : rem);
    return even / A * A;
}
extern "C" void rs_slab_bounds(int d1, int nranks, int rank, int *j0, int *j1)
{
    *j0 = slab_bound(d1, nranks, rank);
    *j1 = slab_bound(d1, nranks, rank + 1);
}
