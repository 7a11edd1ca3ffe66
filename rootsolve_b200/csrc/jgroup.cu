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
//    phi = S*species + 4*j + 2*k + l (3-D), S*species + 2*j + k (2-D) or
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

#include "solver.cuh"
#include "grid_rows.cuh"

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

__global__ void __launch_bounds__(256) jg_wave_kernel(GridDesc g, int S, int nlevels, unsigned *covered,
                                                      int *__restrict__ group_of, int base, int *__restrict__ stats)
{
    cg::grid_group grid = cg::this_grid();
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;          // 32-bit index arithmetic throughout: columns < 2^31,
    const int nthreads = gridDim.x * blockDim.x;                    // and 64-bit divisions would dominate a step
    const int d1 = g.d[0], d2 = g.d[1], d3 = g.d[2];
    const bool three = (g.type == RS_TYPE_3D), one = (g.type == RS_TYPE_1D);
    // candidate lines: (species, j, k) in 3-D, (species, j) in 2-D, (species) in 1-D; the last coordinate
    // follows from the level
    const int nlines = three ? g.nspec * d1 * d2 : one ? g.nspec : g.nspec * d1;
    const int dl = three ? d3 : one ? d1 : d2;
    int count = 0, top = -1;
    const int nsteps = nlevels + WAVE_SLOTS - 1;
    for (int step = 0; step < nsteps; step++) {
        const int slot_lo = max(0, step - nlevels + 1), slot_hi = min(WAVE_SLOTS - 1, step);
        const int work = (slot_hi - slot_lo + 1) * nlines;
        for (int wq = tid; wq < work; wq += nthreads) {
            const int slot = slot_lo + (int)(wq / nlines);
            const int q = wq % nlines;
            const int t = step - slot;
            int s, last;
            int c;
            if (three) {
                const int k = (int)(q % d2);
                const int r = q / d2;
                const int j = (int)(r % d1);
                s = (int)(r / d1);
                last = t - S * s - 4 * j - 2 * k;
                c = q * d3 + last;
            } else if (one) {
                s = (int)q;
                last = t - S * s;
                c = q * d1 + last;
            } else {
                const int j = (int)(q % d1);
                s = (int)(q / d1);
                last = t - S * s - 2 * j;
                c = q * d2 + last;
            }
            if (last < 0 || last >= dl) continue;
            if (group_of[c] >= 0) continue;          // taken by an earlier group (earlier step or pass)
            const unsigned bit = 1u << slot;
            bool clash = false;
            emit_column(g, (int)c + 1, [&](int row) { clash = clash || (covered[row - 1] & bit); });
            if (clash) continue;
            group_of[c] = base + slot;
            emit_column(g, (int)c + 1, [&](int row) { atomicOr(&covered[row - 1], bit); });
            count++;
            top = max(top, base + slot);
        }
        grid.sync();
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

int jgroup_device(Ctx &c, const GridDesc &g, int n, int nnz, const int *d_ian, const int *d_jan,
                  const int *d_rowptr, const int *d_col, int maxg, DevBuf<int> &group_of, int *ngp)
{
    (void)nnz;
    RS_TRY(group_of.alloc(n));
    DevBuf<unsigned char> state;
    DevBuf<int> counters;  // [0] undecided, [1] accepted
    RS_TRY(state.alloc(n));
    RS_TRY(counters.alloc(2));
    RS_TRY(state.zero(c.stream));
    int total = 0, groups = 0;

    if (wave_applicable(g)) {
        DevBuf<unsigned> covered;   // one bit per group slot of the current pass
        RS_TRY(covered.alloc(n));
        const bool three = g.type == RS_TYPE_3D;
        const bool one = g.type == RS_TYPE_1D;
        // species weight S: larger than the biggest level difference between neighbouring cells
        const int S = three ? 5 : one ? 2 : 3;
        const int nlevels = three ? S * (g.nspec - 1) + 4 * (g.d[0] - 1) + 2 * (g.d[1] - 1) + g.d[2]
                            : one ? S * (g.nspec - 1) + g.d[0]
                                  : S * (g.nspec - 1) + 2 * (g.d[0] - 1) + g.d[1];
        int dev = 0, sms = 0, per_sm = 0;
        RS_CUDA(cudaGetDevice(&dev));
        RS_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        RS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, jg_wave_kernel, 256, 0));
        const long long nlines = three ? (long long)g.nspec * g.d[0] * g.d[1] : one ? (long long)g.nspec : (long long)g.nspec * g.d[0];
        long long want = (nlines * WAVE_SLOTS + 255) / 256;
        const char *ctas_env = getenv("RS_WAVE_CTAS");
        const int ctas = ctas_env ? atoi(ctas_env) : 4;
        long long cap = (long long)sms * (per_sm > ctas ? ctas : per_sm);
        int blocks = (int)(want < cap ? want : cap);
        if (blocks < 1) blocks = 1;
        fill_int_kernel<<<grid_for(n, 256), 256, 0, c.stream>>>(n, group_of.p, -1);
        c.launches++;
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
            int *go = group_of.p, *st = counters.p;
            void *args[] = {&gg, &Sx, &nl, &cv, &go, &bs, &st};
            {
                Launch L(c, K_JGROUP, (double)WAVE_SLOTS * 4.0 * n);
                RS_CUDA(cudaLaunchCooperativeKernel((void *)jg_wave_kernel, dim3(blocks), dim3(256), args, 0, c.stream));
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
