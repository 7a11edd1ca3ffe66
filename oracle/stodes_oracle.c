/*
 * stodes_oracle.c -- CPU oracle (TEST INFRASTRUCTURE ONLY, see stodes_oracle.h).
 *
 * Restates, in C, the algorithm of rootSolve's stodes path.  Each function
 * cites the reference lines it follows.  Arrays handed to the Yale routines are
 * addressed 1-based through pointers shifted by one element, because the
 * compressed-pointer bookkeeping of the package is written around 1..n ranges
 * with 0 as the list terminator.
 */
#include "stodes_oracle.h"

#include <math.h>
#include <setjmp.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* ------------------------------------------------------------------------ */
/* Structure of the Jacobian for 1-D / 2-D / 3-D multi-species grids          */
/* ------------------------------------------------------------------------ */

/* src/sparsecoeff.f:5-83.  Column K=(species i, box j): self, next box (or the
 * wrap), previous box (or the wrap), then the other species of the box. */
int orc_sparse1d(int ntot, int nspec, int dim, int cyclic, int *nnz, int *ian, int *jan)
{
    (void)dim;
    if ((ntot / nspec) * nspec != ntot) return -1;
    const int nb = ntot / nspec, cap = *nnz;
    int ij = 1;
    ian[0] = 1;
    for (int i = 1; i <= nspec; i++) {
        const int isp = (i - 1) * nb;
        for (int j = 1; j <= nb; j++) {
            const int K = isp + j;
            int rows[3], nr = 0;
            rows[nr++] = K;
            if (j < nb) rows[nr++] = K + 1;
            else if (cyclic == 1) rows[nr++] = isp + 1;
            if (j > 1) rows[nr++] = K - 1;
            else if (cyclic == 1) rows[nr++] = isp + nb;
            if (ij - 1 + nr + (nspec - 1) > cap) return -2;
            for (int q = 0; q < nr; q++) jan[ij++ - 1] = rows[q];
            for (int L = 1; L <= nspec; L++)
                if (L != i) jan[ij++ - 1] = (L - 1) * nb + j;
            ian[K] = ij;
        }
    }
    *nnz = ij - 1;
    return 0;
}

/* src/sparsecoeff.f:89-176.  Column M=(i,j,k): self, k+1, j+1, j-1, k-1, other
 * species.  No de-duplication; the k==1 cyclic wrap uses j*dimens(1)
 * (line 159) -- kept as is. */
int orc_sparse2d(int ntot, int nspec, const int *dimens, const int *cyclic, int *nnz, int *ian, int *jan)
{
    if ((ntot / nspec) * nspec != ntot) return -1;
    const int d1 = dimens[0], d2 = dimens[1], nb = d1 * d2, cap = *nnz;
    int ij = 1;
    ian[0] = 1;
    for (int i = 1; i <= nspec; i++) {
        const int isp = (i - 1) * nb;
        for (int j = 1; j <= d1; j++)
            for (int k = 1; k <= d2; k++) {
                const int M = isp + (j - 1) * d2 + k;
                int rows[5], nr = 0;
                rows[nr++] = M;
                if (k < d2) rows[nr++] = M + 1;
                else if (cyclic[1] == 1) rows[nr++] = isp + (j - 1) * d2 + 1;
                if (j < d1) rows[nr++] = M + d2;
                else if (cyclic[0] == 1) rows[nr++] = isp + k;
                if (j > 1) rows[nr++] = M - d2;
                else if (cyclic[0] == 1) rows[nr++] = isp + (d1 - 1) * d2 + k;
                if (k > 1) rows[nr++] = M - 1;
                else if (cyclic[1] == 1) rows[nr++] = isp + j * d1;
                if (ij - 1 + nr + (nspec - 1) > cap) return -2;
                for (int q = 0; q < nr; q++) jan[ij++ - 1] = rows[q];
                for (int L = 1; L <= nspec; L++)
                    if (L != i) jan[ij++ - 1] = (L - 1) * nb + (j - 1) * d2 + k;
                ian[M] = ij;
            }
    }
    *nnz = ij - 1;
    return 0;
}

/* src/sparsecoeff.f:178-201: append ival to the current column unless present */
static int put_unique(int *ij, int cap, const int *ian, int *jan, int M, int ival)
{
    for (int q = ian[M - 1]; q <= *ij - 1; q++)
        if (jan[q - 1] == ival) return 0;
    if (*ij > cap) return -2;
    jan[*ij - 1] = ival;
    (*ij)++;
    return 0;
}

/* src/sparsecoeff.f:207-312.  Column M=(i,j,k,l): self, l+1, l-1, k+1, k-1, j+1,
 * j-1, other species; wraps only for cyclic dimensions longer than 2. */
int orc_sparse3d(int ntot, int nspec, const int *dimens, const int *cyclic, int *nnz, int *ian, int *jan)
{
    if ((ntot / nspec) * nspec != ntot) return -1;
    const int d1 = dimens[0], d2 = dimens[1], d3 = dimens[2];
    const int nb = d1 * d2 * d3, cap = *nnz, s1 = d2 * d3;
    int ij = 1, rc = 0;
    ian[0] = 1;
    for (int i = 1; i <= nspec; i++) {
        const int isp = (i - 1) * nb;
        for (int j = 1; j <= d1; j++)
            for (int k = 1; k <= d2; k++)
                for (int l = 1; l <= d3; l++) {
                    const int line = isp + (j - 1) * s1 + (k - 1) * d3;
                    const int M = line + l;
                    rc |= put_unique(&ij, cap, ian, jan, M, M);
                    if (l < d3) rc |= put_unique(&ij, cap, ian, jan, M, M + 1);
                    else if (cyclic[2] == 1 && d3 > 2) rc |= put_unique(&ij, cap, ian, jan, M, line + 1);
                    if (l > 1) rc |= put_unique(&ij, cap, ian, jan, M, M - 1);
                    else if (cyclic[2] == 1 && d3 > 2) rc |= put_unique(&ij, cap, ian, jan, M, line + d3);
                    if (k < d2) rc |= put_unique(&ij, cap, ian, jan, M, M + d3);
                    else if (cyclic[1] == 1 && d2 > 2) rc |= put_unique(&ij, cap, ian, jan, M, isp + (j - 1) * s1 + l);
                    if (k > 1) rc |= put_unique(&ij, cap, ian, jan, M, M - d3);
                    else if (cyclic[1] == 1 && d2 > 2) rc |= put_unique(&ij, cap, ian, jan, M, isp + j * s1 - d3 + l);
                    if (j < d1) rc |= put_unique(&ij, cap, ian, jan, M, M + s1);
                    else if (cyclic[0] == 1 && d1 > 2) rc |= put_unique(&ij, cap, ian, jan, M, isp + (k - 1) * d3 + l);
                    if (j > 1) rc |= put_unique(&ij, cap, ian, jan, M, M - s1);
                    else if (cyclic[0] == 1 && d1 > 2) rc |= put_unique(&ij, cap, ian, jan, M, isp + (d1 - 1) * s1 + (k - 1) * d3 + l);
                    for (int L = 1; L <= nspec; L++)
                        if (L != i) rc |= put_unique(&ij, cap, ian, jan, M, (L - 1) * nb + (j - 1) * s1 + (k - 1) * d3 + l);
                    if (rc) return -2;
                    ian[M] = ij;
                }
    }
    *nnz = ij - 1;
    return 0;
}

/* src/testMap.f:5-23 (updatejan): append the NEW number of full-grid state ii when it is present.  The
 * reference errors as soon as the write position passes the capacity, i.e. already when the capacity is used up
 * exactly.  An index outside 1..totn (only the 2-D k==1 wrap of line 97 can produce one, when d1 > d2) reads
 * beyond `pres` in the reference; here it counts as absent. */
static int put_mapped(int *ij, int cap, int *jan, const int *pres, int totn, int ii)
{
    if (ii < 1 || ii > totn || pres[ii - 1] <= 0) return 0;
    if (*ij > cap) return -2;           /* already failed: callers OR the codes and return afterwards */
    jan[*ij - 1] = pres[ii - 1];
    (*ij)++;
    return (*ij > cap) ? -2 : 0;
}

/* src/testMap.f:25-115.  As sparse2d, on the states present in the map: pres[M-1] = new number (1..ntot, in
 * grid order) of full-grid state M, <= 0 if absent.  Same entry order as sparse2d (incl. the line-97 wrap). */
int orc_sparse2dmap(int ntot, int nspec, const int *dimens, const int *cyclic, int *nnz, int *ian, int *jan,
                    const int *pres, int totn)
{
    if ((ntot / nspec) * nspec != ntot) return -1;
    const int d1 = dimens[0], d2 = dimens[1], nb = d1 * d2, cap = *nnz;
    int ij = 1, rc = 0;
    ian[0] = 1;
    for (int i = 1; i <= nspec; i++) {
        const int isp = (i - 1) * nb;
        for (int j = 1; j <= d1; j++)
            for (int k = 1; k <= d2; k++) {
                const int M = isp + (j - 1) * d2 + k;
                const int Mnew = pres[M - 1];
                if (Mnew <= 0) continue;
                rc |= put_mapped(&ij, cap, jan, pres, totn, M);
                if (k < d2) rc |= put_mapped(&ij, cap, jan, pres, totn, M + 1);
                else if (cyclic[1] == 1) rc |= put_mapped(&ij, cap, jan, pres, totn, isp + (j - 1) * d2 + 1);
                if (j < d1) rc |= put_mapped(&ij, cap, jan, pres, totn, M + d2);
                else if (cyclic[0] == 1) rc |= put_mapped(&ij, cap, jan, pres, totn, isp + k);
                if (j > 1) rc |= put_mapped(&ij, cap, jan, pres, totn, M - d2);
                else if (cyclic[0] == 1) rc |= put_mapped(&ij, cap, jan, pres, totn, isp + (d1 - 1) * d2 + k);
                if (k > 1) rc |= put_mapped(&ij, cap, jan, pres, totn, M - 1);
                else if (cyclic[1] == 1) rc |= put_mapped(&ij, cap, jan, pres, totn, isp + j * d1);
                for (int L = 1; L <= nspec; L++)
                    if (L != i) rc |= put_mapped(&ij, cap, jan, pres, totn, (L - 1) * nb + (j - 1) * d2 + k);
                if (rc) return -2;
                ian[Mnew] = ij;
            }
    }
    *nnz = ij - 1;
    return 0;
}

/* src/testMap.f:119-227.  As sparse3d on the present states; no `interact` search here (the reference appends
 * blindly; with wraps only for dimensions > 2 no duplicate can arise). */
int orc_sparse3dmap(int ntot, int nspec, const int *dimens, const int *cyclic, int *nnz, int *ian, int *jan,
                    const int *pres, int totn)
{
    if ((ntot / nspec) * nspec != ntot) return -1;
    const int d1 = dimens[0], d2 = dimens[1], d3 = dimens[2];
    const int nb = d1 * d2 * d3, cap = *nnz, s1 = d2 * d3;
    int ij = 1, rc = 0;
    ian[0] = 1;
    for (int i = 1; i <= nspec; i++) {
        const int isp = (i - 1) * nb;
        for (int j = 1; j <= d1; j++)
            for (int k = 1; k <= d2; k++)
                for (int l = 1; l <= d3; l++) {
                    const int line = isp + (j - 1) * s1 + (k - 1) * d3;
                    const int M = line + l;
                    const int Mnew = pres[M - 1];
                    if (Mnew <= 0) continue;
                    rc |= put_mapped(&ij, cap, jan, pres, totn, M);
                    if (l < d3) rc |= put_mapped(&ij, cap, jan, pres, totn, M + 1);
                    else if (cyclic[2] == 1 && d3 > 2) rc |= put_mapped(&ij, cap, jan, pres, totn, line + 1);
                    if (l > 1) rc |= put_mapped(&ij, cap, jan, pres, totn, M - 1);
                    else if (cyclic[2] == 1 && d3 > 2) rc |= put_mapped(&ij, cap, jan, pres, totn, line + d3);
                    if (k < d2) rc |= put_mapped(&ij, cap, jan, pres, totn, M + d3);
                    else if (cyclic[1] == 1 && d2 > 2) rc |= put_mapped(&ij, cap, jan, pres, totn, isp + (j - 1) * s1 + l);
                    if (k > 1) rc |= put_mapped(&ij, cap, jan, pres, totn, M - d3);
                    else if (cyclic[1] == 1 && d2 > 2) rc |= put_mapped(&ij, cap, jan, pres, totn, isp + j * s1 - d3 + l);
                    if (j < d1) rc |= put_mapped(&ij, cap, jan, pres, totn, M + s1);
                    else if (cyclic[0] == 1 && d1 > 2) rc |= put_mapped(&ij, cap, jan, pres, totn, isp + (k - 1) * d3 + l);
                    if (j > 1) rc |= put_mapped(&ij, cap, jan, pres, totn, M - s1);
                    else if (cyclic[0] == 1 && d1 > 2)
                        rc |= put_mapped(&ij, cap, jan, pres, totn, isp + (d1 - 1) * s1 + (k - 1) * d3 + l);
                    for (int L = 1; L <= nspec; L++)
                        if (L != i) rc |= put_mapped(&ij, cap, jan, pres, totn, (L - 1) * nb + (j - 1) * s1 + (k - 1) * d3 + l);
                    if (rc) return -2;
                    ian[Mnew] = ij;
                }
    }
    *nnz = ij - 1;
    return 0;
}

/* ------------------------------------------------------------------------ */
/* Column grouping                                                            */
/* ------------------------------------------------------------------------ */

/* src/sparse.f:531-616: greedy first fit.  Group ng takes, scanning columns in
 * ascending order, every not-yet-grouped column none of whose rows is already
 * covered by the group.  Stops at the first empty group. */
int orc_jgroup(int n, const int *ia, const int *ja, int maxg, int *ngrp, int *igp, int *jgp)
{
    char *covered = (char *)malloc((size_t)n + 1);
    char *done = (char *)calloc((size_t)n + 1, 1);
    int ncol = 1, toomuch = 0, ng, rc = 0;
    for (ng = 1; ng <= n; ng++) {
        int start = ncol;
        if (ng <= maxg) igp[ng - 1] = ncol; else toomuch = 1;
        memset(covered, 0, (size_t)n + 1);
        for (int j = 1; j <= n; j++) {
            if (done[j]) continue;
            const int k0 = ia[j - 1], k1 = ia[j] - 1;
            int clash = 0;
            for (int k = k0; k <= k1; k++)
                if (covered[ja[k - 1]]) { clash = 1; break; }
            if (clash) continue;
            jgp[ncol - 1] = j;
            ncol++;
            done[j] = 1;
            for (int k = k0; k <= k1; k++) covered[ja[k - 1]] = 1;
        }
        if (ncol == start) break;     /* empty group: grouping complete */
    }
    if (ng > n) {                     /* loop ran out without an empty group */
        if (ncol <= n) rc = 1;
        ng = maxg;
    }
    *ngrp = ng - 1;
    if (rc == 0 && toomuch) rc = 2;
    free(covered);
    free(done);
    return rc;
}

/* ------------------------------------------------------------------------ */
/* Small numerics                                                             */
/* ------------------------------------------------------------------------ */

/* src/steady.f:208-226 */
double orc_perturb(double *value)
{
    double d = fabs(*value) * 1.0e-8;
    if (d < 1.0e-8) d = 1.0e-8;
    *value = *value + d;
    return d;
}

/* src/sparse.f:264-293 */
void orc_errset(int n, int itol, const double *rtol, const double *atol, const double *y, double *ewt)
{
    for (int i = 0; i < n; i++) {
        const double rt = (itol == 3 || itol == 4) ? rtol[i] : rtol[0];
        const double at = (itol == 2 || itol == 4) ? atol[i] : atol[0];
        ewt[i] = rt * fabs(y[i]) + at;
    }
}

/* src/sparse.f:461-525 */
void orc_sparse_jacob(orc_deriv_t f, int n, const int *ian, const int *jan, const int *igp,
                      const int *jgp, int ngp, double *svar, double *dsvar, double *beta,
                      double t, double *out, int *ipar, double *a)
{
    double *copy = (double *)malloc(sizeof(double) * (size_t)n);
    int neq = n;
    f(&neq, &t, svar, dsvar, out, ipar);
    for (int i = 0; i < n; i++) { beta[i] = -dsvar[i]; copy[i] = svar[i]; }
    int jmin = igp[0];
    for (int ng = 1; ng <= ngp; ng++) {
        const int jmax = igp[ng] - 1;
        for (int j = jmin; j <= jmax; j++) orc_perturb(&svar[jgp[j - 1] - 1]);
        f(&neq, &t, svar, dsvar, out, ipar);
        for (int j = jmin; j <= jmax; j++) {
            const int jj = jgp[j - 1];
            const double divdelt = svar[jj - 1] - copy[jj - 1];   /* the rounded increment */
            svar[jj - 1] = copy[jj - 1];
            for (int k = ian[jj - 1]; k <= ian[jj] - 1; k++) {
                const int i = jan[k - 1];
                a[k - 1] = (dsvar[i - 1] + beta[i - 1]) / divdelt;
            }
        }
        jmin = jmax + 1;
    }
    free(copy);
}

/* ------------------------------------------------------------------------ */
/* Yale sparse matrix package: minimum degree ordering                        */
/* (all arrays 1-based below this line)                                       */
/* ------------------------------------------------------------------------ */

/* src/sparse.f:810-889: build the adjacency lists of M + M' and the degree lists */
static int md_init(int n, const int *ia, const int *ja, int max, int *v, int *l, int *head,
                   int *last, int *next, int *mark, int tag)
{
    for (int vi = 1; vi <= n; vi++) { mark[vi] = 1; l[vi] = 0; head[vi] = 0; }
    int sfs = n + 1;
    for (int vi = 1; vi <= n; vi++) {
        for (int j = ia[vi]; j <= ia[vi + 1] - 1; j++) {
            const int vj = ja[j];
            if (vj == vi) continue;
            if (vj < vi) {               /* strict lower triangle: was a(vj,vi) entered already? */
                int lvk = vi, seen = 0;
                for (int k = 1; k <= mark[vi] - 1; k++) {
                    lvk = l[lvk];
                    if (v[lvk] == vj) { seen = 1; break; }
                }
                if (seen) continue;
            }
            if (sfs >= max) return 9 * n + vi;
            mark[vi]++; v[sfs] = vj; l[sfs] = l[vi]; l[vi] = sfs; sfs++;
            mark[vj]++; v[sfs] = vi; l[sfs] = l[vj]; l[vj] = sfs; sfs++;
        }
    }
    for (int vi = 1; vi <= n; vi++) {
        const int dvi = mark[vi];
        next[vi] = head[dvi];
        head[dvi] = vi;
        last[vi] = -dvi;
        if (next[vi] > 0) last[next[vi]] = vi;
        mark[vi] = tag;
    }
    return 0;
}

/* src/sparse.f:893-947: form element ek from the uneliminated neighbours of vk */
static void md_form_element(int vk, int *tail, int *v, int *l, int *last, int *next, int *mark)
{
    const int tag = mark[vk];
    int s, ls = l[vk];
    *tail = vk;
    while ((s = ls) != 0) {
        ls = l[s];
        const int vs = v[s];
        if (next[vs] >= 0) {             /* uneliminated vertex */
            mark[vs] = tag;
            l[*tail] = s;
            *tail = s;
        } else {                         /* active element: absorb its boundary */
            int lb = l[vs];
            const int nb = last[vs];
            for (int q = 1; q <= nb; q++) {
                const int b = lb;
                lb = l[b];
                const int vb = v[b];
                if (mark[vb] < tag) { mark[vb] = tag; l[*tail] = b; *tail = b; }
            }
            mark[vs] = tag;              /* now inactive */
        }
    }
    l[*tail] = 0;
}

/* src/sparse.f:951-1040: purge inactive elements, mass elimination */
static void md_purge(int *k, int ek, int *tail, int *v, int *l, int *head, int *last, int *next, int *mark)
{
    const int tag = mark[ek];
    int freeslot = 1;
    int li = ek;
    const int cnt = last[ek];
    for (int q = 1; q <= cnt; q++) {
        const int i = li;
        li = l[i];
        const int vi = v[li];
        if (last[vi] != 0) {             /* unlink vi from its degree list */
            if (last[vi] > 0) next[last[vi]] = next[vi];
            else head[-last[vi]] = next[vi];
            if (next[vi] > 0) last[next[vi]] = last[vi];
        }
        int ls = vi;                     /* drop inactive elements from vi's element list */
        for (;;) {
            const int s = ls;
            ls = l[s];
            if (ls == 0) break;
            if (mark[v[ls]] >= tag) { freeslot = ls; l[s] = l[ls]; ls = s; }
        }
        const int lvi = l[vi];
        if (lvi == 0) {                  /* interior vertex: eliminate it now */
            l[i] = l[li];
            li = i;
            (*k)++;
            next[vi] = -(*k);
            last[ek]--;
            continue;
        }
        int classified = 0;
        if (l[lvi] == 0) {
            const int evi = v[lvi];
            if (next[evi] < 0) {
                if (mark[evi] < 0) {     /* duplicate vertex */
                    last[vi] = 0;
                    mark[evi]--;
                } else {                 /* prototype vertex: move to end of boundary list */
                    last[vi] = evi;
                    mark[evi] = -1;
                    l[*tail] = li;
                    *tail = li;
                    l[i] = l[li];
                    li = i;
                }
                classified = 1;
            }
        }
        if (!classified) last[vi] = -ek; /* degree must be recomputed */
        v[freeslot] = ek;                /* insert ek in element list of vi */
        l[freeslot] = l[vi];
        l[vi] = freeslot;
    }
    l[*tail] = 0;
}

/* src/sparse.f:1041-1133: update degrees of the uneliminated vertices in ek */
static void md_update(int ek, int *dmin, int *v, int *l, int *head, int *last, int *next, int *mark)
{
    int tag = mark[ek] - last[ek];
    int i = ek;
    const int cnt = last[ek];
    for (int q = 1; q <= cnt; q++) {
        i = l[i];
        const int vi = v[i];
        int dvi;
        if (last[vi] == 0) continue;
        if (last[vi] < 0) {              /* merge elements to compute the degree */
            tag++;
            dvi = last[ek];
            int s = l[vi], outmatched = 0, es = 0;
            for (;;) {
                s = l[s];
                if (s == 0) break;
                es = v[s];
                if (next[es] >= 0) { mark[es] = tag; dvi++; continue; }
                if (mark[es] < 0) { outmatched = 1; break; }
                int b = es;
                const int nb = last[es];
                for (int p = 1; p <= nb; p++) {
                    b = l[b];
                    const int vb = v[b];
                    if (mark[vb] < tag) { mark[vb] = tag; dvi++; }
                }
            }
            if (outmatched) {            /* adjust overlaps, no degree */
                last[vi] = 0;
                mark[es]--;
                for (;;) {
                    s = l[s];
                    if (s == 0) break;
                    es = v[s];
                    if (mark[es] < 0) mark[es]--;
                }
                continue;
            }
        } else {                         /* prototype: inclusion/exclusion */
            const int evi = last[vi];
            dvi = last[ek] + last[evi] + mark[evi];
            mark[evi] = 0;
        }
        next[vi] = head[dvi];
        head[dvi] = vi;
        last[vi] = -dvi;
        if (next[vi] > 0) last[next[vi]] = vi;
        if (dvi < *dmin) *dmin = dvi;
    }
}

/* src/sparse.f:666-806 */
static int md_order(int n, const int *ia, const int *ja, int max, int *v, int *l, int *head,
                    int *last, int *next, int *mark)
{
    int tag = 0;
    int flag = md_init(n, ia, ja, max, v, l, head, last, next, mark, tag);
    if (flag) return flag;
    int k = 0, dmin = 1, tail = 0;
    while (k < n) {
        while (head[dmin] <= 0) dmin++;
        const int vk = head[dmin];
        head[dmin] = next[vk];
        if (head[dmin] > 0) last[head[dmin]] = -dmin;
        k++;
        next[vk] = -k;
        last[vk] = dmin - 1;
        tag += last[vk];
        mark[vk] = tag;
        md_form_element(vk, &tail, v, l, last, next, mark);
        md_purge(&k, vk, &tail, v, l, head, last, next, mark);
        md_update(vk, &dmin, v, l, head, last, next, mark);
    }
    for (int q = 1; q <= n; q++) {
        next[q] = -next[q];
        last[next[q]] = q;
    }
    return 0;
}

/* src/sparse.f:1245-1423, path 1 only (the only path dsparse uses, line 101) */
int orc_odrv(int n, const int *ia, const int *ja, int *p, int *ip, int nsp, int *isp, int path)
{
    if (path != 1) return 11 * n + 1;
    const int max = (nsp - n) / 2;
    if (max < n) return 10 * n + 1;
    int *v = isp - 1;            /* isp(1..max)        */
    int *l = isp + max - 1;      /* isp(max+1..2max)   */
    int *head = isp + 2 * max - 1;
    /* mark shares storage with v, as in the reference call */
    return md_order(n, ia - 1, ja - 1, max, v, l, head, p - 1, ip - 1, v);
}

/* ------------------------------------------------------------------------ */
/* Yale sparse matrix package: nonsymmetric LDU, compressed pointer storage   */
/* ------------------------------------------------------------------------ */

/* src/sparse.f:1701-1936 (code 1892-1935): sort every row of (ja,a) by the new
 * column order, in place. */
static int ys_reorder_rows(int n, const int *ic, const int *ia, int *ja, double *a, int *jar,
                           double *ar, int *p)
{
    for (int k = 1; k <= n; k++) {
        const int jmin = ia[k], jmax = ia[k + 1] - 1;
        if (jmin > jmax) continue;
        p[n + 1] = n + 1;
        for (int j = jmin; j <= jmax; j++) {
            const int newj = ic[ja[j]];
            int i = n + 1;
            while (p[i] < newj) i = p[i];
            if (p[i] == newj) return n + k;      /* duplicate entry */
            p[newj] = p[i];
            p[i] = newj;
            jar[newj] = ja[j];
            ar[newj] = a[j];
        }
        int i = n + 1;
        for (int j = jmin; j <= jmax; j++) {
            i = p[i];
            ja[j] = jar[i];
            a[j] = ar[i];
        }
    }
    return 0;
}

/* insert vj into the sorted linked list q (head q[np1]); returns 1 if new,
 * 0 if already present.  *qm is the running cursor when merging sorted runs. */
static int q_insert_from(int *q, int start, int vj, int *where)
{
    int m, qm = start;
    do { m = qm; qm = q[m]; } while (qm < vj);
    if (qm == vj) { *where = qm; return 0; }
    q[m] = vj;
    q[vj] = qm;
    *where = vj;
    return 1;
}

/* src/sparse.f:1937-2283: symbolic LDU factorisation */
static int ys_symbolic(int n, int *r, const int *ic, const int *ia, const int *ja, int jlmax, int *il,
                       int *jl, int *ijl, int jumax, int *iu, int *ju, int *iju, int *q, int *ira,
                       int *jra, int *irac, int *irl, int *jrl, int *iru, int *jru)
{
    const int np1 = n + 1;
    int jlmin = 1, jlptr = 0, jumin = 1, juptr = 0;
    il[1] = 1;
    iu[1] = 1;
    for (int k = 1; k <= n; k++) { irac[k] = 0; jra[k] = 0; jrl[k] = 0; jru[k] = 0; }
    /* column pointers for a */
    for (int k = 1; k <= n; k++) {
        const int rk = r[k], iak = ia[rk];
        if (iak >= ia[rk + 1]) return n + rk;              /* null row */
        const int jaiak = ic[ja[iak]];
        if (jaiak > k) return 5 * n + k;                   /* null pivot */
        jra[k] = irac[jaiak];
        irac[jaiak] = k;
        ira[k] = iak;
    }
    for (int k = 1; k <= n; k++) {
        int luk, lastid, lasti, qm, where;
        /* ---- k-th column of L ---- */
        q[np1] = np1;
        luk = -1;
        for (int vj = irac[k]; vj != 0; vj = jra[vj]) {
            if (!q_insert_from(q, np1, vj, &where)) return 2 * n + r[k];  /* duplicate */
            luk++;
        }
        lastid = 0; lasti = 0;
        ijl[k] = jlptr;
        for (int i = jru[k]; i != 0; i = jru[i]) {
            const int jmin = irl[i], jmax = ijl[i] + il[i + 1] - il[i] - 1;
            int lng = jmax - jmin;
            if (lng < 0) continue;
            const int jtmp = jl[jmin];
            if (jtmp != k) lng++;
            if (jtmp == k) r[i] = -r[i];
            if (lastid < lng) { lasti = i; lastid = lng; }
            qm = np1;
            for (int j = jmin; j <= jmax; j++) {          /* merge column i into column k */
                if (q_insert_from(q, qm, jl[j], &where)) luk++;
                qm = where;
            }
        }
        qm = q[np1];
        if (qm != k) return 5 * n + k;
        if (luk != 0) {
            int stored = 0;
            if (lastid == luk) {                          /* identical to tail of column lasti */
                const int irll = irl[lasti];
                ijl[k] = irll + 1;
                if (jl[irll] != k) ijl[k]--;
                stored = 1;
            } else if (jlmin <= jlptr) {                  /* try to overlap the previous column */
                qm = q[qm];
                int j;
                for (j = jlmin; j <= jlptr; j++) {
                    if (jl[j] < qm) continue;
                    break;
                }
                if (j <= jlptr && jl[j] == qm) {
                    ijl[k] = j;
                    int i, mismatch = 0;
                    for (i = j; i <= jlptr; i++) {
                        if (jl[i] != qm) { mismatch = 1; break; }
                        qm = q[qm];
                        if (qm > n) { stored = 1; break; }
                    }
                    if (!stored && !mismatch) jlptr = j - 1;
                }
            }
            if (!stored) {
                jlmin = jlptr + 1;
                ijl[k] = jlmin;
                jlptr += luk;
                if (jlptr > jlmax) return 3 * n + k;
                qm = q[np1];
                for (int j = jlmin; j <= jlptr; j++) { qm = q[qm]; jl[j] = qm; }
            }
        }
        irl[k] = ijl[k];
        il[k + 1] = il[k] + luk;

        /* ---- k-th row of U ---- */
        q[np1] = np1;
        luk = -1;
        {
            const int rk = r[k];
            for (int j = ira[k]; j <= ia[rk + 1] - 1; j++) {
                if (!q_insert_from(q, np1, ic[ja[j]], &where)) return 2 * n + rk;
                luk++;
            }
        }
        lastid = 0; lasti = 0;
        iju[k] = juptr;
        for (int i = jrl[k], i1; i != 0; i = i1) {
            i1 = jrl[i];
            const int jmin = iru[i], jmax = iju[i] + iu[i + 1] - iu[i] - 1;
            int lng = jmax - jmin;
            if (lng < 0) continue;
            const int jtmp = ju[jmin];
            if (jtmp != k) {                              /* advance irl/jrl of column i */
                lng++;
                const int cend = ijl[i] + il[i + 1] - il[i];
                irl[i]++;
                if (irl[i] < cend) {
                    const int j = jl[irl[i]];
                    jrl[i] = jrl[j];
                    jrl[j] = i;
                }
            }
            if (lastid < lng) { lasti = i; lastid = lng; }
            qm = np1;
            for (int j = jmin; j <= jmax; j++) {          /* merge row i into row k */
                if (q_insert_from(q, qm, ju[j], &where)) luk++;
                qm = where;
            }
        }
        if (il[k + 1] > il[k]) {
            const int j = jl[irl[k]];
            jrl[k] = jrl[j];
            jrl[j] = k;
        }
        qm = q[np1];
        if (qm != k) return 5 * n + k;
        if (luk != 0) {
            int stored = 0;
            if (lastid == luk) {
                const int irul = iru[lasti];
                iju[k] = irul + 1;
                if (ju[irul] != k) iju[k]--;
                stored = 1;
            } else if (jumin <= juptr) {
                qm = q[qm];
                int j;
                for (j = jumin; j <= juptr; j++) {
                    if (ju[j] < qm) continue;
                    break;
                }
                if (j <= juptr && ju[j] == qm) {
                    iju[k] = j;
                    int i, mismatch = 0;
                    for (i = j; i <= juptr; i++) {
                        if (ju[i] != qm) { mismatch = 1; break; }
                        qm = q[qm];
                        if (qm > n) { stored = 1; break; }
                    }
                    if (!stored && !mismatch) juptr = j - 1;
                }
            }
            if (!stored) {
                jumin = juptr + 1;
                iju[k] = jumin;
                juptr += luk;
                if (juptr > jumax) return 6 * n + k;
                qm = q[np1];
                for (int j = jumin; j <= juptr; j++) { qm = q[qm]; ju[j] = qm; }
            }
        }
        iru[k] = iju[k];
        iu[k + 1] = iu[k] + luk;

        /* ---- update iru, jru ---- */
        for (int i = k;;) {
            const int i1 = jru[i];
            if (r[i] < 0) {
                r[i] = -r[i];
            } else {
                const int rend = iju[i] + iu[i + 1] - iu[i];
                if (iru[i] < rend) {
                    const int j = ju[iru[i]];
                    jru[i] = jru[j];
                    jru[j] = i;
                }
            }
            i = i1;
            if (i == 0) break;
            iru[i]++;
        }
        /* ---- update ira, jra, irac ---- */
        for (int i = irac[k], i1; i != 0; i = i1) {
            i1 = jra[i];
            ira[i]++;
            if (ira[i] >= ia[r[i] + 1]) continue;
            const int jairai = ic[ja[ira[i]]];
            if (jairai > i) continue;
            jra[i] = irac[jairai];
            irac[jairai] = i;
        }
    }
    ijl[n] = jlptr;
    iju[n] = juptr;
    return 0;
}

/* src/sparse.f:2287-2447: numeric LDU + solve M x = b */
static int ys_numeric(int n, const int *r, const int *c, const int *ic, const int *ia, const int *ja,
                      const double *a, double *z, const double *b, int lmax, const int *il,
                      const int *jl, const int *ijl, double *l, double *d, int umax, const int *iu,
                      const int *ju, const int *iju, double *u, double *row, double *tmp, int *irl,
                      int *jrl)
{
    if (il[n + 1] - 1 > lmax) return 4 * n + 1;
    if (iu[n + 1] - 1 > umax) return 7 * n + 1;
    for (int k = 1; k <= n; k++) { irl[k] = il[k]; jrl[k] = 0; }
    for (int k = 1; k <= n; k++) {
        /* reverse jrl and clear row where the k-th row of L will fill in */
        row[k] = 0.0;
        int i1 = 0;
        for (int i = jrl[k], i2; i != 0; i = i2) {
            i2 = jrl[i];
            jrl[i] = i1;
            i1 = i;
            row[i] = 0.0;
        }
        {
            const int jmin = iju[k], jmax = jmin + iu[k + 1] - iu[k] - 1;
            for (int j = jmin; j <= jmax; j++) row[ju[j]] = 0.0;
        }
        const int rk = r[k];
        for (int j = ia[rk]; j <= ia[rk + 1] - 1; j++) row[ic[ja[j]]] = a[j];
        double sum = b[rk];
        for (int i = i1; i != 0; i = jrl[i]) {
            const double lki = -row[i];
            l[irl[i]] = -lki;
            sum = sum + lki * tmp[i];
            const int jmin = iu[i], jmax = iu[i + 1] - 1;
            if (jmin <= jmax) {
                const int mu = iju[i] - jmin;
                for (int j = jmin; j <= jmax; j++) row[ju[mu + j]] = row[ju[mu + j]] + lki * u[j];
            }
        }
        if (row[k] == 0.0) return 8 * n + k;
        const double dk = 1.0 / row[k];
        d[k] = dk;
        tmp[k] = sum * dk;
        if (k == n) continue;
        {
            const int jmin = iu[k], jmax = iu[k + 1] - 1;
            if (jmin <= jmax) {
                const int mu = iju[k] - jmin;
                for (int j = jmin; j <= jmax; j++) u[j] = row[ju[mu + j]] * dk;
            }
        }
        /* update irl and jrl, keeping jrl in decreasing order */
        for (int i = i1, nxt; i != 0; i = nxt) {
            irl[i]++;
            nxt = jrl[i];
            if (irl[i] >= il[i + 1]) continue;
            const int ijlb = irl[i] - il[i] + ijl[i];
            int j = jl[ijlb];
            while (i <= jrl[j]) j = jrl[j];
            jrl[i] = jrl[j];
            jrl[j] = i;
        }
        if (irl[k] < il[k + 1]) {
            const int j = jl[ijl[k]];
            jrl[k] = jrl[j];
            jrl[j] = k;
        }
    }
    /* back substitution U x = tmp */
    for (int k = n; k >= 1; k--) {
        double sum = tmp[k];
        const int jmin = iu[k], jmax = iu[k + 1] - 1;
        if (jmin <= jmax) {
            const int mu = iju[k] - jmin;
            for (int j = jmin; j <= jmax; j++) sum = sum - u[j] * tmp[ju[mu + j]];
        }
        tmp[k] = sum;
        z[c[k]] = sum;
    }
    return 0;
}

/* src/sparse.f:2507-2563: solve M' x = b with the LDU factors */
static void ys_solve_transposed(int n, const int *r, const int *c, const int *il, const int *jl,
                                const int *ijl, const double *l, const double *d, const int *iu,
                                const int *ju, const int *iju, const double *u, double *z,
                                const double *b, double *tmp)
{
    for (int k = 1; k <= n; k++) tmp[k] = b[c[k]];
    for (int k = 1; k <= n; k++) {           /* U' y = b, forward */
        const int jmin = iu[k], jmax = iu[k + 1] - 1;
        const double tmpk = -tmp[k];
        if (jmin > jmax) continue;
        const int mu = iju[k] - jmin;
        for (int j = jmin; j <= jmax; j++) tmp[ju[mu + j]] = tmp[ju[mu + j]] + tmpk * u[j];
    }
    for (int k = n; k >= 1; k--) {           /* L' x = y, backward */
        double sum = -tmp[k];
        const int jmin = il[k], jmax = il[k + 1] - 1;
        if (jmin <= jmax) {
            const int ml = ijl[k] - jmin;
            for (int j = jmin; j <= jmax; j++) sum = sum + l[j] * tmp[jl[ml + j]];
        }
        tmp[k] = -sum * d[k];
        z[r[k]] = tmp[k];
    }
}

/* src/sparse.f:1427-1697: storage carving + dispatch (paths 1, 2, 4 as used by
 * dsparse).  isp and rsp are separate arrays here, as call_stsparse.c:122-126
 * allocates them (rsp[nsp] doubles, isp[2*nsp] ints). */
int orc_cdrv(int n, int *r0, int *c0, int *ic0, int *ia0, int *ja0, double *a0, double *b0, double *z0,
             int nsp, int *isp0, double *rsp0, int *esp, int path)
{
    const int lratio = 2;
    int *r = r0 - 1, *c = c0 - 1, *ic = ic0 - 1, *ia = ia0 - 1, *ja = ja0 - 1, *isp = isp0 - 1;
    double *a = a0 - 1, *b = b0 - 1, *z = z0 - 1, *rsp = rsp0 - 1;
    if (path < 1 || path > 5) return 11 * n + 1;
    const int il = 1, ijl = il + (n + 1), iu = ijl + n, iju = iu + (n + 1), irl = iju + n,
              jrl = irl + n, jl = jrl + n;
    int flag;
    if (path == 1 || path == 5) {
        const int max = (lratio * nsp + 1 - jl) - (n + 1) - 5 * n;
        int jlmax = max / 2;
        const int q = jl + jlmax, ira = q + (n + 1), jra = ira + n, irac = jra + n, iru = irac + n,
                  jru = iru + n, jutmp = jru + n;
        int jumax = lratio * nsp + 1 - jutmp;
        *esp = max / lratio;
        if (jlmax <= 0 || jumax <= 0) return 10 * n + 1;
        int reordered = 0;
        for (int i = 1; i <= n; i++) if (c[i] != i) { reordered = 1; break; }
        if (reordered) {
            const int ar = nsp + 1 - n;
            flag = ys_reorder_rows(n, ic, ia, ja, a, isp + il - 1, rsp + ar - 1, isp + iu - 1);
            if (flag) return flag;
        }
        flag = ys_symbolic(n, r, ic, ia, ja, jlmax, isp + il - 1, isp + jl - 1, isp + ijl - 1, jumax,
                           isp + iu - 1, isp + jutmp - 1, isp + iju - 1, isp + q - 1, isp + ira - 1,
                           isp + jra - 1, isp + irac - 1, isp + irl - 1, isp + jrl - 1, isp + iru - 1,
                           isp + jru - 1);
        if (flag) return flag;
        /* move ju next to jl */
        jlmax = isp[ijl + n - 1];
        const int ju = jl + jlmax;
        jumax = isp[iju + n - 1];
        for (int j = 1; j <= jumax; j++) isp[ju + j - 1] = isp[jutmp + j - 1];
    }
    const int jlmax = isp[ijl + n - 1];
    const int ju = jl + jlmax;
    const int jumax = isp[iju + n - 1];
    const int l = (ju + jumax - 2 + lratio) / lratio + 1;
    const int lmax = isp[il + n] - 1;
    const int d = l + lmax;
    const int u = d + n;
    const int row = nsp + 1 - n;
    const int tmp = row - n;
    const int umax = tmp - u;
    *esp = umax - (isp[iu + n] - 1);
    if (path == 1 || path == 2) {
        if (umax < 0) return 10 * n + 1;
        flag = ys_numeric(n, r, c, ic, ia, ja, a, z, b, lmax, isp + il - 1, isp + jl - 1, isp + ijl - 1,
                          rsp + l - 1, rsp + d - 1, umax, isp + iu - 1, isp + ju - 1, isp + iju - 1,
                          rsp + u - 1, rsp + row - 1, rsp + tmp - 1, isp + irl - 1, isp + jrl - 1);
        if (flag) return flag;
    }
    if (path == 4)
        ys_solve_transposed(n, r, c, isp + il - 1, isp + jl - 1, isp + ijl - 1, rsp + l - 1, rsp + d - 1,
                            isp + iu - 1, isp + ju - 1, isp + iju - 1, rsp + u - 1, z, b, rsp + tmp - 1);
    if (path == 3) return 11 * n + 1;   /* nnsc: never used by dsparse, not restated */
    return 0;
}

/* ------------------------------------------------------------------------ */
/* Driver                                                                     */
/* ------------------------------------------------------------------------ */

static const char *yale_message(int flag, int n)
{
    switch (flag / n) {                 /* src/sparse.f:207-258 */
    case 1: return "sparse solver: null row in a";
    case 2: return "sparse solver: duplicate entry in a";
    case 3: case 6: return "insufficient storage in nsfc - increase argument lrw";
    case 4: case 7: return "insufficient storage in nnfc - increase lrw";
    case 5: return "sparse solver: null pivot";
    case 8: return "sparse solver: zero pivot";
    case 9: return "insufficient storage in md - increase lrw";
    case 10: return "insufficient storage in cdrv/odrv-increase lrw";
    default: return "illegal path specifications";
    }
}

static int fail(orc_result *res, int code, const char *msg)
{
    res->status = code;
    snprintf(res->msg, sizeof res->msg, "%s", msg);
    return code;
}

/* src/call_stsparse.c:60-320 (allocation, ipar/out packing, result packing) +
 * src/sparse.f:30-195 (the Newton loop) + src/sparse.f:300-455 (dispatch). */
int orc_stsparse(const orc_args *A, orc_result *R)
{
    const int n = A->neq, nsp = A->lrw, maxg = A->ngp, maxit = A->maxiter;
    const int type = A->type, atype = abs(type);
    int nnz = A->nnzvec[0];
    const double t0 = now_s();
    R->status = 0; R->msg[0] = 0; R->niter = 0; R->steady = 0;
    R->t_struct = R->t_order = R->t_jac = R->t_factor = R->t_solve = 0.0;

    /* ipar / out, src/steady_utils.c:90-126 (DLL branch) */
    const int lrpar = A->nout + A->nrpar, lipar = 3 + A->nipar;
    double *out = (double *)calloc((size_t)(lrpar > 0 ? lrpar : 1), sizeof(double));
    int *ipar = (int *)calloc((size_t)lipar, sizeof(int));
    ipar[0] = A->nout; ipar[1] = lrpar; ipar[2] = lipar;
    for (int j = 0; j < A->nipar; j++) ipar[3 + j] = A->ipar[j];
    for (int j = 0; j < A->nrpar; j++) out[A->nout + j] = A->rpar[j];

    int *Rp = (int *)calloc((size_t)n, sizeof(int)), *Cp = (int *)calloc((size_t)n, sizeof(int)),
        *ICp = (int *)calloc((size_t)n, sizeof(int));
    double *rsp = (double *)calloc((size_t)nsp, sizeof(double));
    int *isp = (int *)calloc((size_t)2 * nsp, sizeof(int));
    double *svar = (double *)malloc(sizeof(double) * n), *dsvar = (double *)calloc(n, sizeof(double)),
           *beta = (double *)calloc(n, sizeof(double)), *x = (double *)calloc(n, sizeof(double)),
           *ewt = (double *)calloc(n, sizeof(double));
    double *alpha = (double *)calloc((size_t)(nnz > 0 ? nnz : 1), sizeof(double));
    int *ian = R->ian, *jan = R->jan, *igp = R->igp, *jgp = R->jgp;
    int dims[7] = {0, 0, 0, 0, 0, 0, 0};
    int rc = 0, ngp = 0, esp = 0, neq = n;
    double tin = A->time;
    memcpy(svar, R->y, sizeof(double) * n);       /* caller put y0 in R->y */
    memset(ian, 0, sizeof(int) * ((size_t)n + 1));
    memset(jan, 0, sizeof(int) * (size_t)nnz);
    if (type == 0) {
        memcpy(ian, A->ian_in, sizeof(int) * ((size_t)n + 1));
        memcpy(jan, A->jan_in, sizeof(int) * (size_t)nnz);
    }
    { /* src/call_stsparse.c:189-195 */
        const int nd = (atype == 2) ? 3 : (atype == 3 || atype == 30) ? 5 : (atype == 4 || atype == 40) ? 7 : 0;
        for (int j = 0; j < nd && j + 1 < A->nnnz; j++) dims[j] = A->nnzvec[j + 1];
    }
    /* mapped grids, src/call_stsparse.c:196-206: after the cyclic flags come TotN and the map itself */
    const int *pres = NULL;
    int totn = 0;
    if (atype == 30 || atype == 40) {
        const int at = (atype == 30) ? 6 : 8;
        if (A->nnnz < at + 1 || A->nnnz < at + 1 + A->nnzvec[at]) { rc = fail(R, -11, "nnz vector too short for a mapped grid"); goto done; }
        totn = A->nnzvec[at];
        pres = A->nnzvec + at + 1;
    }
    memset(igp, 0, sizeof(int) * ((size_t)maxg + 1));
    memset(jgp, 0, sizeof(int) * (size_t)n);
    for (int j = 0; j < maxit; j++) R->precis[j] = 0.0;

    const int positivity = A->posit > 0;
    orc_errset(n, A->itol, A->rtol, A->atol, svar, ewt);

    /* ---- xsparsestruct, src/sparse.f:300-455 ---- */
    double ts = now_s();
    if (atype == 1) {
        A->func(&neq, &tin, svar, beta, out, ipar);
        ian[0] = 1;
        int ij = 0, enough = 1;
        for (int i = 0; i < n; i++) {
            const double keep = svar[i];
            const double divdelt = orc_perturb(&svar[i]);
            A->func(&neq, &tin, svar, dsvar, out, ipar);
            for (int j = 0; j < n; j++)
                if (fabs((dsvar[j] - beta[j]) / divdelt) > 1e-30) {
                    ij++;
                    if (ij > nnz) enough = 0;
                    if (enough) jan[ij - 1] = j + 1;
                }
            ian[i + 1] = ij + 1;
            svar[i] = keep;
        }
        if (!enough) { rc = fail(R, -10, "error during determining sparsity: nnz too small"); goto done; }
        nnz = ij;
    } else if (atype == 2) {
        rc = orc_sparse1d(n, dims[0], dims[1], dims[2], &nnz, ian, jan);
    } else if (atype == 3) {
        rc = orc_sparse2d(n, dims[0], &dims[1], &dims[3], &nnz, ian, jan);
    } else if (atype == 4) {
        rc = orc_sparse3d(n, dims[0], &dims[1], &dims[4], &nnz, ian, jan);
    } else if (atype == 30) {
        rc = orc_sparse2dmap(n, dims[0], &dims[1], &dims[3], &nnz, ian, jan, pres, totn);
    } else if (atype == 40) {
        rc = orc_sparse3dmap(n, dims[0], &dims[1], &dims[4], &nnz, ian, jan, pres, totn);
    } else if (atype != 0) {
        rc = fail(R, -11, "cannot run stodes: sparsetype not known"); goto done;
    }
    if (rc == -1) { fail(R, rc, "cannot generate sparse jacobian - N and nspec not compatible"); goto done; }
    if (rc == -2) { fail(R, rc, "cannot generate sparse jacobian - not enough room for nonzeros"); goto done; }
    if (type >= 0) {
        const int ier = orc_jgroup(n, ian, jan, maxg, &ngp, igp, jgp);
        if (ier == 1) { rc = fail(R, -12, "not enough memory for JGROUP"); goto done; }
        if (ier == 2) { rc = fail(R, -13, "error during grouping: NGP too small"); goto done; }
        /* Reference quirk: with one column per group (ngp == n) JGROUP's loop ends before igp(ngp+1) is written
         * (src/sparse.f:569-600), it stays 0, and xSparseJacob then skips the last group: the last column of the
         * Jacobian is never filled.  quirk_unterminated_igp = 0 repairs the terminator instead. */
        if (!A->quirk_unterminated_igp && ngp == n && igp[ngp] == 0) igp[ngp] = n + 1;
    }
    dims[0] = nnz;
    dims[1] = ngp;
    R->t_struct = now_s() - ts;
    if (type < 0) goto pack;            /* sparsereturn */

    /* ---- ordering, src/sparse.f:101-105 ---- */
    ts = now_s();
    rc = orc_odrv(n, ian, jan, Rp, ICp, nsp, isp, 1);
    if (rc) { fail(R, rc, yale_message(rc, n)); goto done; }
    for (int k = 1; k <= n; k++) Cp[ICp[k - 1] - 1] = k;
    R->t_order = now_s() - ts;

    {
        int path = A->quirk_first_step_transposed ? 2 : 4;
        for (int j = 0; j < n; j++) x[j] = 0.0;
        for (int it = 1; it <= maxit; it++) {
            R->niter = it;
            ts = now_s();
            orc_sparse_jacob(A->func, n, ian, jan, igp, jgp, ngp, svar, dsvar, beta, tin, out, ipar, alpha);
            R->t_jac += now_s() - ts;
            if (it == 1) {
                if (R->a0) memcpy(R->a0, alpha, sizeof(double) * (size_t)nnz);
                if (R->beta0) memcpy(R->beta0, beta, sizeof(double) * (size_t)n);
            }
            /* convergence test 1, src/sparse.f:122-133 */
            double prec = 0.0, maxewt = 0.0;
            for (int k = 0; k < n; k++) {
                prec += fabs(beta[k]);
                const double w = fabs(beta[k] / ewt[k]);
                if (w > maxewt) maxewt = w;
            }
            R->precis[it - 1] = prec / n;
            if (maxewt <= 1.0) { R->steady = 1; break; }

            ts = now_s();
            rc = orc_cdrv(n, Rp, Cp, ICp, ian, jan, alpha, beta, x, nsp, isp, rsp, &esp, 1);
            if (rc) { fail(R, rc, yale_message(rc, n)); goto done; }
            R->t_factor += now_s() - ts;
            ts = now_s();
            rc = orc_cdrv(n, Rp, Cp, ICp, ian, jan, alpha, beta, x, nsp, isp, rsp, &esp, path);
            if (rc) { fail(R, rc, yale_message(rc, n)); goto done; }
            R->t_solve += now_s() - ts;
            path = 4;

            /* update, src/sparse.f:152-163 */
            double relchange = 0.0;
            for (int k = 0; k < n; k++) {
                if (fabs(x[k]) > relchange) relchange = fabs(x[k]);
                svar[k] = svar[k] + x[k];
                if (positivity && svar[k] < 0.0) svar[k] = 0.0;
            }
            if (!positivity && A->ipos > 1)
                for (int k = 0; k < A->ipos; k++)
                    if (svar[A->pos[k] - 1] < 0.0) svar[A->pos[k] - 1] = 0.0;
            /* convergence test 2, src/sparse.f:165-183 */
            if (relchange <= A->ctol) {
                if (it < maxit) {
                    for (int j = 0; j < n; j++) beta[j] = 0.0;
                    A->func(&neq, &tin, svar, beta, out, ipar);
                    double s = 0.0;
                    for (int j = 0; j < n; j++) s += fabs(beta[j]);
                    R->precis[it] = s;              /* a sum, not a mean (line 177) */
                    R->niter = it + 1;
                }
                R->steady = 1;
                break;
            }
            orc_errset(n, A->itol, A->rtol, A->atol, svar, ewt);
        }
        dims[2] = nsp - esp;
    }

pack:
    /* src/call_stsparse.c:278-298 */
    memcpy(R->y, svar, sizeof(double) * n);
    if (A->nout > 0) {
        A->func(&neq, &tin, svar, dsvar, out, ipar);
        for (int j = 0; j < A->nout; j++) R->y[n + j] = out[j];
    }
    R->dims[0] = dims[0]; R->dims[1] = dims[1]; R->dims[2] = dims[2];
    if (A->imethod == 10) { R->steady = 0; R->niter = 1; R->precis[0] = 0.0; }
done:
    R->t_total = now_s() - t0;
    free(out); free(ipar); free(Rp); free(Cp); free(ICp); free(rsp); free(isp);
    free(svar); free(dsvar); free(beta); free(x); free(ewt); free(alpha);
    return R->status;
}
