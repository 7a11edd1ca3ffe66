// grid_rows.cuh -- the rows of one Jacobian column on a regular multi-species grid, in the order the
// reference emits them (sparse1d / sparse2d / sparse3d, src/sparsecoeff.f:5-312).  Shared by the structure
// generator (structure.cu) and the wavefront column grouping (jgroup.cu).
#pragma once

#include "solver.cuh"

namespace rs {

// Emits the 1-based row indices of column M (1-based) in the reference's order
// into sink(row).  The stencil part is de-duplicated for 3-D only, as `interact`
// does (src/sparsecoeff.f:178-201); 1-D and 2-D append blindly.
template <typename Sink>
__device__ __forceinline__ void emit_column(const GridDesc &g, int M, Sink &&sink)
{
    const int nb = g.ncell;
    const int i = (M - 1) / nb;            // species, 0-based
    const int cell = (M - 1) - i * nb;     // 0-based cell
    const int isp = i * nb;
    if (g.type == RS_TYPE_1D) {
        const int j = cell + 1, N = nb;
        sink(M);
        if (j < N) sink(M + 1); else if (g.cyc[0] == 1) sink(isp + 1);
        if (j > 1) sink(M - 1); else if (g.cyc[0] == 1) sink(isp + N);
    } else if (g.type == RS_TYPE_2D) {
        const int d1 = g.d[0], d2 = g.d[1];
        const int j = cell / d2 + 1, k = cell - (j - 1) * d2 + 1;
        sink(M);
        if (k < d2) sink(M + 1); else if (g.cyc[1] == 1) sink(isp + (j - 1) * d2 + 1);
        if (j < d1) sink(M + d2); else if (g.cyc[0] == 1) sink(isp + k);
        if (j > 1) sink(M - d2); else if (g.cyc[0] == 1) sink(isp + (d1 - 1) * d2 + k);
        if (k > 1) sink(M - 1); else if (g.cyc[1] == 1) sink(isp + j * d1);  // sic: src/sparsecoeff.f:159
    } else {
        const int d1 = g.d[0], d2 = g.d[1], d3 = g.d[2], s1 = d2 * d3;
        const int j = cell / s1 + 1;
        const int rem = cell - (j - 1) * s1;
        const int k = rem / d3 + 1, l = rem - (k - 1) * d3 + 1;
        const int line = isp + (j - 1) * s1 + (k - 1) * d3;
        int seen[7], ns = 0;
        auto put = [&](int v) {
            for (int q = 0; q < ns; q++) if (seen[q] == v) return;
            seen[ns++] = v;
            sink(v);
        };
        put(M);
        if (l < d3) put(M + 1); else if (g.cyc[2] == 1 && d3 > 2) put(line + 1);
        if (l > 1) put(M - 1); else if (g.cyc[2] == 1 && d3 > 2) put(line + d3);
        if (k < d2) put(M + d3); else if (g.cyc[1] == 1 && d2 > 2) put(isp + (j - 1) * s1 + l);
        if (k > 1) put(M - d3); else if (g.cyc[1] == 1 && d2 > 2) put(isp + j * s1 - d3 + l);
        if (j < d1) put(M + s1); else if (g.cyc[0] == 1 && d1 > 2) put(isp + (k - 1) * d3 + l);
        if (j > 1) put(M - s1); else if (g.cyc[0] == 1 && d1 > 2) put(isp + (d1 - 1) * s1 + (k - 1) * d3 + l);
    }
    // other species of the same box, ascending (a different species can never
    // duplicate a stencil entry, so no search is needed here)
    for (int L = 0; L < g.nspec; L++)
        if (L != i) sink(L * nb + cell + 1);
}


}  // namespace rs
