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
    if (rs::g_comm) { ncclCommDestroy(rs::g_comm); rs::g_comm = nullptr; rs::g_nranks = 1; rs::g_rank = 0; }
#endif
}

// Planes [j0, j1) of the slowest grid dimension owned by `rank`: as even as possible, the first
// d1 % nranks ranks get one more.
extern "C" void rs_slab_bounds(int d1, int nranks, int rank, int *j0, int *j1)
{
    const int base = d1 / nranks, rem = d1 % nranks;
    *j0 = rank * base + (rank < rem ? rank : rem);
    *j1 = *j0 + base + (rank < rem ? 1 : 0);
}
