// grid_stencil.cuh -- the row of the Jacobian on a non-cyclic regular grid with one or two species, as a pure
// function of the row's grid position: candidate columns in ascending order (the order of the row-wise arrays).
// Shared by the implicit-column SpMV (krylov.cu) and the implicit-column FD assembly (fdjac.cu); krylov_setup
// verifies the pattern row by row (grid_cols_check_kernel) before either is used.
#pragma once

namespace rs {

struct GridCols { int d1, d2, d3, s1, ncell; };

template <int NSPEC>
struct GridRow {
    static constexpr int NSLOT = NSPEC == 1 ? 7 : 9;
    int c[NSLOT];
    bool has[NSLOT];
    // candidate columns of row r in ascending order; has[q]: the entry exists
    __device__ __forceinline__ void set(const GridCols &g, int r, bool valid)
    {
        int sp = 0, cell = r;
        if (NSPEC == 2 && r >= g.ncell) { sp = 1; cell = r - g.ncell; }
        const int j = cell / g.s1, rem = cell - j * g.s1, k = rem / g.d3, l = rem - k * g.d3;
        constexpr int o = NSPEC == 1 ? 0 : 1;
        if (NSPEC == 2) {
            has[0] = valid && sp == 1; c[0] = cell;                    // species 0 of the same cell
            has[NSLOT - 1] = valid && sp == 0; c[NSLOT - 1] = cell + g.ncell;   // species 1 of the same cell
        }
        has[o + 0] = valid && j > 0;        c[o + 0] = r - g.s1;
        has[o + 1] = valid && k > 0;        c[o + 1] = r - g.d3;
        has[o + 2] = valid && l > 0;        c[o + 2] = r - 1;
        has[o + 3] = valid;                 c[o + 3] = r;
        has[o + 4] = valid && l < g.d3 - 1; c[o + 4] = r + 1;
        has[o + 5] = valid && k < g.d2 - 1; c[o + 5] = r + g.d3;
        has[o + 6] = valid && j < g.d1 - 1; c[o + 6] = r + g.s1;
    }
};

}  // namespace rs
