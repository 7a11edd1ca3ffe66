// jgroup.cu -- K1: column grouping for the finite-difference Jacobian.
//
// Reproduces JGROUP (src/sparse.f:531-616) bit for bit.  The reference is a
// sequential greedy: group g takes, scanning columns in ascending order, every
// ungrouped column none of whose rows is covered by a column already taken
// into g.  Equivalently: column j joins g iff no ungrouped column j' < j that
// shares a row with j has joined g -- the lexicographically first maximal
// independent set of the column-intersection graph.  Two exact device variants:
//
//  * wavefront (regular non-cyclic 1-D/2-D/3-D grids): with the level function
//    phi = S*species + 3*j + 2*k + l (3-D), S*species + 2*j + k (2-D) or
//    S*species + j (1-D) every pair of conflicting columns lies on different
//    levels and the lower index is always on the lower level, so sweeping the
//    levels in order with a "row covered" bitmap is the sequential algorithm,
//    run one hyperplane at a time by a cooperative kernel (grid-wide barrier
//    between levels) and pipelined over 16 groups at once (see jg_wave_kernel);
//  * rounds (any structure, used for cyclic and user patterns): each
//    undecided column looks at the lower-index columns it conflicts with;
//    one of them in the group -> rejected; all of them rejected or grouped
//    earlier -> accepted; otherwise it waits for the next round.  Decisions are
//    only ever taken on final information, so the fixed point is the greedy's.
#include <cooperative_groups.h>

#include <cstdlib>
#include <cstring>

#include "solver.cuh"
#include "grid_rows.cuh"

#ifndef RS_WAVE_DEFAULT_VARIANT
#define RS_WAVE_DEFAULT_VARIANT 0   // index into the variant table of jgroup_device
#endif

namespace cg = cooperative_groups;

namespace rs {

enum : unsigned char { ST_UNDECIDED = 0, ST_ACCEPTED = 1, ST_REJECTED = 2, ST_DONE = 3 };

__global__ void jg_round_kernel(int n, const int *__restrict__ ian, const int *__restrict__ jan,
                                const int *__restrict__ rowptr, const int *__restrict__ col,
                                volatile unsigned char *state, int *__restrict__ undecided)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n || state[j] != ST_UNDECIDED) return;
    for (int attempt = 0; attempt < 8; attempt++) {
        bool pending = false, reject = false;
        for (int k = ian[j] - 1; k < ian[j + 1] - 1 && !reject; k++) {
            const int i = jan[k] - 1;
            for (int p = rowptr[i]; p < rowptr[i + 1]; p++) {
                const int jj = col[p];
                if (jj >= j) break;  // columns of a row are sorted ascending
                const unsigned char st = state[jj];
                if (st == ST_ACCEPTED) { reject = true; break; }
                if (st == ST_UNDECIDED) pending = true;
            }
        }
        if (reject) { state[j] = ST_REJECTED; return; }
        if (!pending) { state[j] = ST_ACCEPTED; return; }
    }
    atomicAdd(undecided, 1);
}

__global__ void jg_commit_kernel(int n, unsigned char *__restrict__ state, int *__restrict__ group_of, int g,
                                 int *__restrict__ accepted)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    int mine = 0;
    if (j < n) {
        const unsigned char st = state[j];
        if (st == ST_ACCEPTED) { group_of[j] = g; state[j] = ST_DONE; mine = 1; }
        else if (st == ST_REJECTED) state[j] = ST_UNDECIDED;
    }
    const unsigned m = __ballot_sync(0xffffffffu, mine);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(accepted, __popc(m));
}

// The wavefront variant, software-pipelined over the groups.  Group g+1's verdict on a column only
// needs (i) group g's verdict on the same column and (ii) its own verdicts on lower levels, so group
// slot g can sweep level (step - g) while slot g+1 sweeps level (step - g - 1): WAVE_SLOTS groups
// advance together, one grid-wide barrier per step, nlevels + WAVE_SLOTS - 1 steps per pass instead
// of nlevels per group.  covered[row] holds one bit per slot.  Rows of a column come from
// emit_column (same order and content as ian/jan), so no structure array is read.
// group_of[c] < 0 on entry marks the columns still to be grouped; a pass assigns base .. base+SLOTS-1.
constexpr int WAVE_SLOTS = 16;

//
// Levels.  Column c = (species s, cell (j,k,l)) may only depend on columns of a lower level, i.e. on every EARLIER
// column sharing a row with it: same species, cell offset a sum of two stencil offsets -- the worst cases
// (-1,+1,0), (-1,0,+1), (0,-1,+1) ask for a > b > 1 in level = a j + b k + l, so 3 j + 2 k + l (2 j + k in 2-D) --
// or an earlier species at a stencil neighbour (level difference at most a), hence the species weight S = a + 1.
// U: (line, slot) pairs a thread takes through each memory round trip together; MINB: CTAs per SM the register
// budget is cut for (more resident threads = fewer pairs per thread, against spills).
// ONE: single species (branch-free rows, see the sweep).
template <int U, int MINB, bool ONE>
__global__ void __launch_bounds__(256, MINB) jg_wave_kernel(GridDesc g, int S, int nlevels, unsigned *covered,
                                                            int *__restrict__ group_of, int base, int *__restrict__ stats,
                                                            int *__restrict__ line_level)
{
    cg::grid_group grid = cg::this_grid();
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;          // 32-bit index arithmetic throughout: columns < 2^31
    const int nthreads = gridDim.x * blockDim.x;
    const int d1 = g.d[0], d2 = g.d[1], d3 = g.d[2];
    const bool three = (g.type == RS_TYPE_3D), one = (g.type == RS_TYPE_1D);
    // candidate lines: (species, j, k) in 3-D, (species, j) in 2-D, (species) in 1-D; the last coordinate
    // follows from the level.  line_level[q] = level of the line's first cell, computed once: the sweep below
    // visits (line, slot) pairs with shifts and one cached load, no division.
    const int nlines = three ? g.nspec * d1 * d2 : one ? g.nspec : g.nspec * d1;
    const int dl = three ? d3 : one ? d1 : d2;
    // the grid as a 3-D box (leading extents 1): what the single-species path works with
    const int box1 = three ? d1 : 1, box2 = three ? d2 : one ? 1 : d1, box3 = dl;
    for (int q = tid; q < nlines; q += nthreads) {
        int lv;
        if (three) { const int k = q % d2, r = q / d2; lv = S * (r / d1) + 3 * (r % d1) + 2 * k; }
        else if (one) lv = S * q;
        else lv = S * (q / d1) + 2 * (q % d1);
        if (ONE) {   // level and "neighbour line exists" flags (k-1, k+1, j-1, j+1) in one word
            const int kf = q % box2, jf = q / box2;
            lv = (lv << 4) | (kf > 0 ? 1 : 0) | (kf < box2 - 1 ? 2 : 0) | (jf > 0 ? 4 : 0) | (jf < box1 - 1 ? 8 : 0);
        }
        line_level[q] = lv;
    }
    grid.sync();
    int count = 0, top = -1;
    const int nsteps = nlevels + WAVE_SLOTS - 1;
    static_assert(WAVE_SLOTS == 16, "the (line, slot) split below uses 4 bits for the slot");
    const int work = nlines * WAVE_SLOTS;
    // A step is latency bound (a few dependent L2 round trips, then the barrier), so a thread takes WAVE_UNROLL
    // pairs through each round trip together: all their group_of loads, then all their covered loads, then the
    // updates -- the loads of a phase are independent and in flight at once.
    for (int step = 0; step < nsteps; step++) {
        if constexpr (ONE) {
            // Single species: the (up to) 7 rows of a cell are computed branch-free -- a neighbour beyond the grid
            // is replaced by the cell itself -- and every load is unconditional (a pair without a cell on this
            // level reads cell 0 and is ignored), so the loads of all U pairs really are in flight together.
            const int s1 = box2 * box3;
            for (int w0 = tid; w0 < work; w0 += nthreads * U) {
                int cc[U], fl[U], ll[U];
                bool ok[U];
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const int wq = min(w0 + u * nthreads, work - 1);
                    const int packed = line_level[wq >> 4];
                    const int last = step - (wq & (WAVE_SLOTS - 1)) - (packed >> 4);
                    ok[u] = (w0 + u * nthreads < work) && last >= 0 && last < dl;
                    cc[u] = ok[u] ? (wq >> 4) * dl + last : 0;
                    ll[u] = ok[u] ? last : 0;
                    fl[u] = ok[u] ? (packed & 15) : 0;
                }
                int taken[U];
#pragma unroll
                for (int u = 0; u < U; u++) taken[u] = group_of[cc[u]];
                int rows[U][6];
                unsigned seen[U];
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const int cu = cc[u];
                    rows[u][0] = ll[u] > 0 ? cu - 1 : cu;
                    rows[u][1] = ll[u] < box3 - 1 ? cu + 1 : cu;
                    rows[u][2] = (fl[u] & 1) ? cu - box3 : cu;
                    rows[u][3] = (fl[u] & 2) ? cu + box3 : cu;
                    rows[u][4] = (fl[u] & 4) ? cu - s1 : cu;
                    rows[u][5] = (fl[u] & 8) ? cu + s1 : cu;
                    seen[u] = covered[cu] | covered[rows[u][0]] | covered[rows[u][1]] | covered[rows[u][2]] | covered[rows[u][3]] |
                              covered[rows[u][4]] | covered[rows[u][5]];
                }
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const int slot = (w0 + u * nthreads) & (WAVE_SLOTS - 1);
                    const unsigned bit = 1u << slot;
                    if (!ok[u] || taken[u] >= 0 || (seen[u] & bit)) continue;
                    group_of[cc[u]] = base + slot;
                    atomicOr(&covered[cc[u]], bit);
#pragma unroll
                    for (int i = 0; i < 6; i++)
                        if (rows[u][i] != cc[u]) atomicOr(&covered[rows[u][i]], bit);
                    count++;
                    top = max(top, base + slot);
                }
            }
        } else {
            for (int w0 = tid; w0 < work; w0 += nthreads * U) {
                int cc[U];
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const int wq = w0 + u * nthreads;
                    cc[u] = -1;
                    if (wq < work) {
                        const int last = step - (wq & (WAVE_SLOTS - 1)) - line_level[wq >> 4];   // slot g sweeps level step - g
                        if (last >= 0 && last < dl) cc[u] = (wq >> 4) * dl + last;
                    }
                }
                int taken[U];
#pragma unroll
                for (int u = 0; u < U; u++) taken[u] = cc[u] >= 0 ? group_of[cc[u]] : 0;   // >= 0: in an earlier group
                unsigned seen[U];
#pragma unroll
                for (int u = 0; u < U; u++) {
                    seen[u] = 0u;
                    if (cc[u] >= 0 && taken[u] < 0) emit_column(g, cc[u] + 1, [&](int row) { seen[u] |= covered[row - 1]; });
                }
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const int slot = (w0 + u * nthreads) & (WAVE_SLOTS - 1);
                    const unsigned bit = 1u << slot;
                    if (cc[u] < 0 || taken[u] >= 0 || (seen[u] & bit)) continue;
                    group_of[cc[u]] = base + slot;
                    emit_column(g, cc[u] + 1, [&](int row) { atomicOr(&covered[row - 1], bit); });
                    count++;
                    top = max(top, base + slot);
                }
            }
        }
        grid.sync();
    }
    if (count) { atomicAdd(&stats[0], count); atomicMax(&stats[1], top); }
}

// 3-D grids with one species: the same wavefront as a pipeline of planes.  A block owns plane j (planes j, j + B, ... when
// there are more planes than resident blocks) and sweeps the in-plane levels mu = 2 k + l in order, WAVE_SLOTS groups
// pipelined as above (slot g is at level step - g), with a block-wide barrier per step instead of a grid-wide one.
// A column of plane j conflicts with earlier columns of planes j - 1 (at k +- 1, l +- 1: in-plane levels up to mu + 2)
// and j - 2 (same k, l), so plane j may run step t once plane j - 1 has finished step t + 2: each plane publishes the
// number of steps it has finished, its successor waits for that count -- the level order 3 j + 2 k + l of the grid-wide
// variant, enforced between neighbours only.  Marks and verdicts live in L2 (atomics, .cg loads): L1 is never
// consulted, so block barriers order them within a plane and the release / acquire pair on the step counter orders
// them between planes.
// Measured at 256^3: 7.2 ms against 12.8 ms for the grid-wide variant (one block per SM with the registers to keep a
// lane's loads in flight together: 8.3 ms, two planes in turn for most blocks; two blocks per SM at 56 registers: 10.2 ms;
// mark words only fetched for columns without a group: 7.2 ms).  Two other layouts were tried and were slower: marks
// stored by (plane, level, k) so that a warp's loads are contiguous (14.6 ms), and no marks at all -- the twelve earlier
// conflicting columns' group numbers read directly, no atomics (18.6 ms).
constexpr int PLANE_WORKERS = 512;                       // 32 lanes for each of the 16 slots
constexpr int PLANE_THREADS = PLANE_WORKERS + 32;        // + one warp that publishes and polls the step counters
constexpr int PLANE_U = 4;                               // cells a lane takes through each memory round trip together

// the spin reads relaxed (an acquire load invalidates the SM's L1 on every turn); one fence after the wait gives the acquire
__device__ __forceinline__ int ld_relaxed_gpu(const int *p)
{
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(int *p, int v)
{
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(PLANE_THREADS, 1) jg_plane_kernel(int d1, int d2, int d3, unsigned *covered, int *group_of, int base,
                                                                   int *stats, int *progress)
{
    const int tid = threadIdx.x;
    const bool worker = tid < PLANE_WORKERS;
    const int slot = tid & (WAVE_SLOTS - 1), lane = tid >> 4;
    constexpr int LANES = PLANE_WORKERS / WAVE_SLOTS;
    const unsigned bit = 1u << slot;
    const int s1 = d2 * d3;
    const int nlev = 2 * (d2 - 1) + d3, nsteps = nlev + WAVE_SLOTS - 1;
    int count = 0, top = -1;
    for (int p = blockIdx.x; p < d1; p += gridDim.x) {
        int seen = 0;   // steps plane p - 1 is known to have finished (service warp)
        for (int t = 0; t < nsteps; t++) {
            if (!worker) {
                // the step about to start needs plane p - 1 through step t + 2; asked for while the workers are still
                // busy with step t - 1, so the round trip is off their path unless the neighbour really is behind
                if (p > 0 && tid == PLANE_WORKERS) {
                    const int need = min(t + 3, nsteps);
                    if (seen < need) {
                        const long long t0 = clock64();
                        while ((seen = ld_relaxed_gpu(progress + p - 1)) < need) {
                            if (clock64() - t0 > 4000000000LL) __trap();   // ~2 s: never on a healthy run
                            __nanosleep(32);
                        }
                        __threadfence();
                    }
                }
            }
            __syncthreads();   // step t - 1 of this plane is complete, plane p - 1 is far enough ahead
            if (!worker) {
                if (tid == PLANE_WORKERS && t > 0) { __threadfence(); st_release_gpu(progress + p, t); }   // t steps finished
                continue;
            }
            const int mu = t - slot;
            if (mu < 0 || mu >= nlev) continue;
            const int kmin = mu - d3 + 1 > 0 ? (mu - d3 + 2) >> 1 : 0, kmax = min(d2 - 1, mu >> 1);
            const int plane0 = p * s1;
            for (int k0 = kmin + lane; k0 <= kmax; k0 += LANES * PLANE_U) {
                int cc[PLANE_U], rows[PLANE_U][6], taken[PLANE_U];
                unsigned mark[PLANE_U];
                bool ok[PLANE_U];
#pragma unroll
                for (int u = 0; u < PLANE_U; u++) {
                    const int k = k0 + u * LANES;
                    ok[u] = k <= kmax;
                    const int kk = ok[u] ? k : kmin, l = mu - 2 * kk;
                    const int c = plane0 + kk * d3 + l;
                    cc[u] = c;
                    rows[u][0] = l > 0 ? c - 1 : c;
                    rows[u][1] = l < d3 - 1 ? c + 1 : c;
                    rows[u][2] = kk > 0 ? c - d3 : c;
                    rows[u][3] = kk < d2 - 1 ? c + d3 : c;
                    rows[u][4] = p > 0 ? c - s1 : c;
                    rows[u][5] = p < d1 - 1 ? c + s1 : c;
                }
#pragma unroll
                for (int u = 0; u < PLANE_U; u++) taken[u] = ok[u] ? __ldcg(group_of + cc[u]) : 0;
                // the step is bound by the scattered loads (one L1 wavefront per lane: the cells of a level are d3 - 2
                // words apart), so the seven mark words are only fetched for columns that still need a group -- under
                // half of all (cell, slot) pairs
#pragma unroll
                for (int u = 0; u < PLANE_U; u++) {
                    mark[u] = 0u;
                    if (taken[u] < 0)
                        mark[u] = __ldcg(covered + cc[u]) | __ldcg(covered + rows[u][0]) | __ldcg(covered + rows[u][1]) | __ldcg(covered + rows[u][2]) |
                                  __ldcg(covered + rows[u][3]) | __ldcg(covered + rows[u][4]) | __ldcg(covered + rows[u][5]);
                }
#pragma unroll
                for (int u = 0; u < PLANE_U; u++) {
                    if (taken[u] >= 0 || (mark[u] & bit)) continue;
                    __stcg(group_of + cc[u], base + slot);
                    atomicOr(&covered[cc[u]], bit);
#pragma unroll
                    for (int i = 0; i < 6; i++)
                        if (rows[u][i] != cc[u]) atomicOr(&covered[rows[u][i]], bit);
                    count++;
                    top = max(top, base + slot);
                }
            }
        }
        __syncthreads();   // the last step of the plane
        if (tid == PLANE_WORKERS) { __threadfence(); st_release_gpu(progress + p, nsteps); }
    }
    if (count) { atomicAdd(&stats[0], count); atomicMax(&stats[1], top); }
}

__global__ void fill_int_kernel(int n, int *__restrict__ p, int v)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) p[i] = v;
}

static bool wave_applicable(const GridDesc &g)
{
    if (g.type != RS_TYPE_1D && g.type != RS_TYPE_2D && g.type != RS_TYPE_3D) return false;
    for (int q = 0; q < 3; q++) if (g.cyc[q]) return false;
    return true;
}

// Long non-cyclic 1-D grids.  The greedy grouping of sparse1d's pattern is periodic along the column: apart from
// nspec - 1 boxes at either end, box j + 3 repeats box j in every species (first-fit on a translation-invariant
// conflict graph; checked against the sequential JGROUP for 1..16 species and lengths up to 200 000 boxes, tests).
// So the wavefront -- one grid barrier per box, 0.7 s at 200 000 boxes -- runs on a column of ~100 boxes with the
// same length modulo 3, and its head, its period and its tail are laid out over the real column.
constexpr int JG_1D_SHORT = 96;
__global__ void jg_expand_1d_kernel(int nspec, int nb, int nshort, int H, int T, const int *__restrict__ gs, int *__restrict__ g)
{
    const long long total = (long long)nspec * nb;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const int s = (int)(q / nb), j = (int)(q - (long long)s * nb);
        int js;
        if (j < H) js = j;
        else if (j >= nb - T) js = nshort - (nb - j);
        else js = H - 3 + (j - H) % 3;
        g[q] = gs[(long long)s * nshort + js];
    }
}

int jgroup_device(Ctx &c, const GridDesc &g, int n, int nnz, const int *d_ian, const int *d_jan,
                  const int *d_rowptr, const int *d_col, int maxg, DevBuf<int> &group_of, int *ngp)
{
    (void)nnz;
    if (g.type == RS_TYPE_1D && !g.cyc[0] && g.d[0] > 4 * JG_1D_SHORT && g.nspec + 8 < JG_1D_SHORT / 3 && !getenv("RS_NO_JG_PERIODIC")) {
        GridDesc gs = g;
        gs.d[0] = JG_1D_SHORT + (g.d[0] - JG_1D_SHORT) % 3;
        gs.ncell = gs.d[0];
        DevBuf<int> short_groups;
        RS_TRY(jgroup_device(c, gs, g.nspec * gs.d[0], 0, nullptr, nullptr, nullptr, nullptr, maxg, short_groups, ngp));
        RS_TRY(group_of.alloc(n));
        Launch L(c, K_JGROUP, 8.0 * n);
        jg_expand_1d_kernel<<<grid_for(n, 256), 256, 0, c.stream>>>(g.nspec, g.d[0], gs.d[0], g.nspec + 8, g.nspec + 8, short_groups.p, group_of.p);
        RS_CUDA(cudaGetLastError());
        RS_CUDA(cudaStreamSynchronize(c.stream));   // short_groups goes out of scope
        return RS_OK;
    }
    RS_TRY(group_of.alloc(n));
    DevBuf<unsigned char> state;
    DevBuf<int> counters;  // [0] undecided, [1] accepted
    RS_TRY(state.alloc(n));
    RS_TRY(counters.alloc(2));
    RS_TRY(state.zero(c.stream));
    int total = 0, groups = 0;

    if (wave_applicable(g)) {
        const bool three = g.type == RS_TYPE_3D;
        // 3-D, one species: planes pipelined block to block (jg_plane_kernel); everything else: grid-wide levels
        const bool planes = three && g.nspec == 1 && g.d[0] > 1 && !getenv("RS_NO_JG_PLANES");
        DevBuf<unsigned> covered;   // one bit per group slot of the current pass
        RS_TRY(covered.alloc((size_t)n));
        const bool one = g.type == RS_TYPE_1D;
        // species weight S: larger than the biggest level difference between neighbouring cells
        const int S = three ? 4 : one ? 2 : 3;
        const int nlevels = three ? S * (g.nspec - 1) + 3 * (g.d[0] - 1) + 2 * (g.d[1] - 1) + g.d[2]
                            : one ? S * (g.nspec - 1) + g.d[0]
                                  : S * (g.nspec - 1) + 2 * (g.d[0] - 1) + g.d[1];
        // kernel variant (RS_WAVE_VARIANT picks another one for tuning sweeps)
        struct Variant { const void *fn; const char *name; };
        static const Variant variants[] = {
            {(const void *)jg_wave_kernel<1, 6, true>, "u1b6"}, {(const void *)jg_wave_kernel<2, 6, true>, "u2b6"},
            {(const void *)jg_wave_kernel<2, 4, true>, "u2b4"}, {(const void *)jg_wave_kernel<4, 4, true>, "u4b4"},
            {(const void *)jg_wave_kernel<4, 3, true>, "u4b3"}, {(const void *)jg_wave_kernel<2, 8, true>, "u2b8"},
            {(const void *)jg_wave_kernel<8, 2, true>, "u8b2"}, {(const void *)jg_wave_kernel<4, 6, true>, "u4b6"}};
        int vi = RS_WAVE_DEFAULT_VARIANT;
        if (const char *ve = getenv("RS_WAVE_VARIANT"))
            for (int q = 0; q < (int)(sizeof(variants) / sizeof(variants[0])); q++)
                if (strcmp(ve, variants[q].name) == 0) vi = q;
        // several species: the general rows (emit_column), one pair at a time
        const void *wave_fn = g.nspec == 1 ? variants[vi].fn : (const void *)jg_wave_kernel<1, 6, false>;
        int dev = 0, sms = 0, per_sm = 0;
        RS_CUDA(cudaGetDevice(&dev));
        RS_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        RS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wave_fn, 256, 0));
        const long long nlines = three ? (long long)g.nspec * g.d[0] * g.d[1] : one ? (long long)g.nspec : (long long)g.nspec * g.d[0];
        DevBuf<int> line_level;
        RS_TRY(line_level.alloc((size_t)nlines));
        long long want = (nlines * WAVE_SLOTS + 255) / 256;
        const char *ctas_env = getenv("RS_WAVE_CTAS");
        const int ctas = ctas_env ? atoi(ctas_env) : 8;   // CTAs per SM taking part (the kernel's occupancy caps it)
        long long cap = (long long)sms * (per_sm > ctas ? ctas : per_sm);
        int blocks = (int)(want < cap ? want : cap);
        if (blocks < 1) blocks = 1;
        fill_int_kernel<<<grid_for(n, 256), 256, 0, c.stream>>>(n, group_of.p, -1);
        c.launches++;
        DevBuf<int> progress;
        int plane_blocks = 0;
        if (planes) {
            int occ = 0;
            RS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void *)jg_plane_kernel, PLANE_THREADS, 0));
            plane_blocks = std::min(g.d[0], sms * std::max(occ, 1));   // all resident (cooperative launch): a block may wait for another
            RS_TRY(progress.alloc((size_t)g.d[0]));
        }
        int base = 0;
        while (total < n) {
            if (base >= maxg) {
                set_error("error during grouping: NGP too small (should be more than %d)", maxg);
                return RS_EGROUP;
            }
            RS_TRY(covered.zero(c.stream));
            const int init[2] = {0, -1};
            RS_CUDA(cudaMemcpyAsync(counters.p, init, 2 * sizeof(int), cudaMemcpyHostToDevice, c.stream));
            GridDesc gg = g;
            int Sx = S, nl = nlevels, bs = base;
            unsigned *cv = covered.p;
            int *go = group_of.p, *st = counters.p, *ll = line_level.p;
            if (planes) {
                RS_TRY(progress.zero(c.stream));
                int e1 = g.d[0], e2 = g.d[1], e3 = g.d[2];
                int *pg = progress.p;
                void *pargs[] = {&e1, &e2, &e3, &cv, &go, &bs, &st, &pg};
                Launch L(c, K_JGROUP, (double)WAVE_SLOTS * 4.0 * n);
                RS_CUDA(cudaLaunchCooperativeKernel((const void *)jg_plane_kernel, dim3(plane_blocks), dim3(PLANE_THREADS), pargs, 0, c.stream));
            } else {
                void *args[] = {&gg, &Sx, &nl, &cv, &go, &bs, &st, &ll};
                Launch L(c, K_JGROUP, (double)WAVE_SLOTS * 4.0 * n);
                RS_CUDA(cudaLaunchCooperativeKernel(wave_fn, dim3(blocks), dim3(256), args, 0, c.stream));
            }
            int st_h[2] = {0, -1};
            RS_CUDA(cudaMemcpyAsync(st_h, counters.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, c.stream));
            RS_CUDA(cudaStreamSynchronize(c.stream));
            if (st_h[0] == 0) break;  // cannot happen while columns remain; guards against a hang
            total += st_h[0];
            groups = st_h[1] + 1;
            base += WAVE_SLOTS;
        }
        if (groups > maxg) {
            set_error("error during grouping: NGP too small (should be at least %d, while it is %d)", groups, maxg);
            return RS_EGROUP;
        }
    } else {
        const int blocks = (n + 255) / 256;
        while (total < n) {
            if (groups >= maxg) {
                set_error("error during grouping: NGP too small (should be more than %d)", maxg);
                return RS_EGROUP;
            }
            int guard = 0;
            for (;;) {
                RS_CUDA(cudaMemsetAsync(counters.p, 0, 2 * sizeof(int), c.stream));
                {
                    Launch L(c, K_JGROUP, 1.0 * n);
                    jg_round_kernel<<<blocks, 256, 0, c.stream>>>(n, d_ian, d_jan, d_rowptr, d_col, state.p, counters.p);
                }
                int und = 0;
                RS_CUDA(cudaMemcpyAsync(&und, counters.p, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
                RS_CUDA(cudaStreamSynchronize(c.stream));
                if (und == 0) break;
                if (++guard > n + 8) { set_error("internal: grouping rounds did not converge"); return RS_EGROUP; }
            }
            {
                Launch L(c, K_JGROUP, 6.0 * n);
                jg_commit_kernel<<<blocks, 256, 0, c.stream>>>(n, state.p, group_of.p, groups, counters.p + 1);
            }
            int acc_h = 0;
            RS_CUDA(cudaMemcpyAsync(&acc_h, counters.p + 1, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
            RS_CUDA(cudaStreamSynchronize(c.stream));
            if (acc_h == 0) break;
            total += acc_h;
            groups++;
        }
    }
    *ngp = groups;
    return RS_OK;
}

// igp/jgp exactly as JGROUP lays them out: jgp lists the columns group by group,
// ascending inside a group; igp(g) is the 1-based start of group g, igp(ngp+1)=n+1.
int jgroup_lists(Ctx &c, int n, int ngp, const int *d_group_of, int *igp_host, int *jgp_host)
{
    std::vector<int> g((size_t)n);
    RS_CUDA(cudaMemcpyAsync(g.data(), d_group_of, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, c.stream));
    RS_CUDA(cudaStreamSynchronize(c.stream));
    std::vector<int> start((size_t)ngp + 1, 0);
    for (int j = 0; j < n; j++) start[g[j] + 1]++;
    for (int q = 0; q < ngp; q++) start[q + 1] += start[q];
    if (igp_host) {
        for (int q = 0; q <= ngp; q++) igp_host[q] = start[q] + 1;
        // Reference quirk: when every column ends up alone in its group (ngp == n) the DO loop of
        // src/sparse.f:569 runs out before it meets the empty group that would write igp(ngp+1),
        // so that slot keeps the zero call_stsparse.c:208-209 put there.  Reported the same way.
        if (ngp == n) igp_host[ngp] = 0;
    }
    if (jgp_host) {
        std::vector<int> cur(start.begin(), start.end() - 1);
        for (int j = 0; j < n; j++) jgp_host[cur[g[j]]++] = j + 1;
    }
    return RS_OK;
}

}  // namespace rs
