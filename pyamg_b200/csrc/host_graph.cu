// Host-side setup helpers that are inherently sequential in the reference and must give the
// IDENTICAL integer result: greedy MIS colouring (pyamg/amg_core/graph.h:171-188 over
// :34-61) used to build the multicolour Gauss-Seidel ordering, and the three-pass standard
// aggregation (pyamg/amg_core/smoothed_aggregation.h:121-221).  Arrays are HOST pointers.
#include <algorithm>
#include <vector>
#include "common.cuh"

using namespace b200amg;

// colour k = lexicographically-first maximal independent set of the vertices still uncoloured
extern "C" int b200amg_vertex_coloring_mis_host(int n, const int *Ap, const int *Aj, int *color)
{
    if (n < 0) return set_err(B200AMG_EINVAL, "vertex_coloring_mis: n < 0");
    // state: color[i] >= 0 final; UNSEEN (-1) still to be coloured; blocked[i] == round marks
    // vertices excluded from the current round's independent set
    std::vector<int> blocked(n > 0 ? n : 1, -1);
    for (int i = 0; i < n; ++i) color[i] = -1;
    int remaining = n, round = 0;
    while (remaining > 0) {
        for (int i = 0; i < n; ++i) {
            if (color[i] != -1 || blocked[i] == round) continue;
            color[i] = round;
            --remaining;
            for (int p = Ap[i]; p < Ap[i + 1]; ++p) {
                const int j = Aj[p];
                if (color[j] == -1) blocked[j] = round;
            }
        }
        ++round;
    }
    return round;   // number of colours
}

// ---- Jones-Plassmann and largest-degree-first colourings (pyamg/amg_core/graph.h:242-335; pyamg/graph.py:84-125
// methods 'JP', 'LDF').  Each round: one serial pass of the weight-based independent-set selection over the
// uncoloured vertices (a vertex joins when no uncoloured neighbour outweighs it, ties to the larger index; its
// uncoloured neighbours sit the round out), then every vertex picked this round takes the smallest colour no
// neighbour holds.  Integer results identical to the reference for the same weights.
namespace {
struct RoundColouring {
    int n;
    const int *Ap, *Aj;
    int *col;                                   // -1 uncoloured, -2 sitting this round out, >= 0 colour
    std::vector<char> seen;

    int pick(int round, const double *w)
    {
        int picked = 0;
        for (int i = 0; i < n; ++i) {
            if (col[i] != -1) continue;
            bool wins = true;
            for (int p = Ap[i]; p < Ap[i + 1] && wins; ++p) {
                const int j = Aj[p];
                if (col[j] == round) { col[i] = -2; wins = false; }
                else if (col[j] == -1 && (w[j] > w[i] || (w[j] == w[i] && j > i))) wins = false;
            }
            if (!wins) continue;
            for (int p = Ap[i]; p < Ap[i + 1]; ++p) if (col[Aj[p]] == -1) col[Aj[p]] = -2;
            col[i] = round;
            ++picked;
        }
        for (int i = 0; i < n; ++i) if (col[i] == -2) col[i] = -1;
        return picked;
    }

    void smallest_free_colour(int round)
    {
        seen.assign((size_t)round + 1, 0);
        for (int i = 0; i < n; ++i) {
            if (col[i] != round) continue;
            std::fill(seen.begin(), seen.end(), 0);
            for (int p = Ap[i]; p < Ap[i + 1]; ++p) {
                const int j = Aj[p];
                if (j != i && col[j] >= 0) seen[col[j]] = 1;
            }
            int c = 0;
            while (c < round && seen[c]) ++c;
            col[i] = c;
        }
    }
};
}  // namespace

// weights[i] = the random number of vertex i (the degree is added inside, like the reference; the caller's array is
// left untouched).  Returns the number of colours.
extern "C" int b200amg_vertex_coloring_jones_plassmann_host(int n, const int *Ap, const int *Aj, int *color,
                                                            const double *weights)
{
    if (n < 0) return set_err(B200AMG_EINVAL, "vertex_coloring_jones_plassmann: n < 0");
    std::vector<double> w(n > 0 ? n : 1);
    for (int i = 0; i < n; ++i) { color[i] = -1; w[i] = weights[i] + (Ap[i + 1] - Ap[i]); }
    RoundColouring rc{n, Ap, Aj, color, {}};
    int done = 0, round = 0, most = -1;
    while (done < n) {
        done += rc.pick(round, w.data());
        rc.smallest_free_colour(round);
        ++round;
    }
    for (int i = 0; i < n; ++i) most = color[i] > most ? color[i] : most;
    return most + 1;
}

// weight of an uncoloured vertex in a round = its random number + its uncoloured neighbours (self excluded)
extern "C" int b200amg_vertex_coloring_ldf_host(int n, const int *Ap, const int *Aj, int *color, const double *weights)
{
    if (n < 0) return set_err(B200AMG_EINVAL, "vertex_coloring_LDF: n < 0");
    std::vector<double> w(n > 0 ? n : 1);
    for (int i = 0; i < n; ++i) color[i] = -1;
    RoundColouring rc{n, Ap, Aj, color, {}};
    int done = 0, round = 0, most = -1;
    while (done < n) {
        for (int i = 0; i < n; ++i) {
            if (color[i] != -1) continue;
            int open = 0;
            for (int p = Ap[i]; p < Ap[i + 1]; ++p) open += (color[Aj[p]] == -1 && Aj[p] != i);
            w[i] = weights[i] + open;
        }
        done += rc.pick(round, w.data());
        rc.smallest_free_colour(round);
        ++round;
    }
    for (int i = 0; i < n; ++i) most = color[i] > most ? color[i] : most;
    return most + 1;
}

// agg[i] in [0, n_agg) or -1 (isolated node); roots[k] = seed node of aggregate k.
// Returns n_agg (>= 0) or a negative error.
extern "C" int b200amg_standard_aggregation_host(int n, const int *Ap, const int *Aj, int *agg, int *roots)
{
    if (n < 0) return set_err(B200AMG_EINVAL, "standard_aggregation: n < 0");
    enum : int { FREE = -2, ISOLATED = -1 };
    std::vector<char> tentative(n > 0 ? n : 1, 0);   // joined in pass 2 (does not attract others)
    for (int i = 0; i < n; ++i) agg[i] = FREE;
    int nagg = 0;
    // pass 1: a node whose neighbours are all free seeds an aggregate with all of them
    for (int i = 0; i < n; ++i) {
        if (agg[i] != FREE) continue;
        bool has_nbr = false, nbr_taken = false;
        for (int p = Ap[i]; p < Ap[i + 1]; ++p) {
            const int j = Aj[p];
            if (j == i) continue;
            has_nbr = true;
            if (agg[j] != FREE) { nbr_taken = true; break; }
        }
        if (!has_nbr) { agg[i] = ISOLATED; continue; }
        if (nbr_taken) continue;
        roots[nagg] = i;
        agg[i] = nagg;
        for (int p = Ap[i]; p < Ap[i + 1]; ++p) agg[Aj[p]] = nagg;
        ++nagg;
    }
    // pass 2: leftover nodes join the aggregate of their first pass-1-aggregated neighbour
    for (int i = 0; i < n; ++i) {
        if (agg[i] != FREE) continue;
        for (int p = Ap[i]; p < Ap[i + 1]; ++p) {
            const int j = Aj[p];
            if (agg[j] >= 0 && !tentative[j]) { agg[i] = agg[j]; tentative[i] = 1; break; }
        }
    }
    // pass 3: what is still free seeds new aggregates with its free neighbours.  The reference
    // renumbers to 0-based IN THE SAME LOOP and tests "x[j] == 0" for free, so an already
    // renumbered member of aggregate 0 with a smaller index is captured as well
    // (smoothed_aggregation.h:196-218); reproduced for identical aggregates.
    for (int i = 0; i < n; ++i) {
        if (agg[i] != FREE) continue;
        roots[nagg] = i;
        agg[i] = nagg;
        for (int p = Ap[i]; p < Ap[i + 1]; ++p) {
            const int j = Aj[p];
            if (agg[j] == FREE || (j < i && agg[j] == 0)) agg[j] = nagg;
        }
        ++nagg;
    }
    return nagg;
}

// Ruge-Stuben first-pass C/F splitting (amg_core.rs_cf_splitting, pyamg/amg_core/ruge_stuben.h:157-297).
// S = strength graph without diagonal, T = S^T, both CSR on the HOST; split[i] = 1 (C) or 0 (F).
//
// The selection order between nodes of equal measure is defined by the reference's bucket array (nodes
// ordered by measure; raising a node moves it to the far end of its bucket and hands that slot to the
// next bucket, lowering moves it to the near end), so the same mechanics are kept -- including that
// bucket starts exist only for measures < n, that a measure >= n-1 is never raised and 0 never lowered.
namespace {
struct MeasureBuckets {
    std::vector<int> measure, start, size, node_at, slot_of;
    explicit MeasureBuckets(int n) : measure(n + 1, 0), start(n + 1, 0), size(n + 1, 0), node_at(n + 1, 0), slot_of(n + 1, 0) {}
    void exchange(int a, int b)
    {
        const int na = node_at[a], nb = node_at[b];
        node_at[a] = nb; slot_of[nb] = a;
        node_at[b] = na; slot_of[na] = b;
    }
    void raise(int k)
    {
        const int m = measure[k], last = start[m] + size[m] - 1;
        exchange(slot_of[k], last);
        --size[m]; ++size[m + 1];
        start[m + 1] = last;
        measure[k] = m + 1;
    }
    void lower(int k)
    {
        const int m = measure[k];
        exchange(slot_of[k], start[m]);
        --size[m]; ++size[m - 1];
        ++start[m];
        start[m - 1] = start[m] - size[m - 1];
        measure[k] = m - 1;
    }
};
}  // namespace

extern "C" int b200amg_rs_cf_splitting_host(int n, const int *Sp, const int *Sj, const int *Tp, const int *Tj, int *split)
{
    if (n < 0) return set_err(B200AMG_EINVAL, "rs_cf_splitting: n < 0");
    enum : int { F_PT = 0, C_PT = 1, OPEN = 2 };
    MeasureBuckets B(n);
    for (int i = 0; i < n; ++i) ++B.size[B.measure[i] = Tp[i + 1] - Tp[i]];
    for (int v = 0, run = 0; v < n; ++v) { B.start[v] = run; run += B.size[v]; B.size[v] = 0; }
    for (int i = 0; i < n; ++i) {
        const int m = B.measure[i], at = B.start[m] + B.size[m]++;
        B.node_at[at] = i;
        B.slot_of[i] = at;
    }
    // nobody depends on the node (or only itself): F point from the start
    for (int i = 0; i < n; ++i) {
        const int m = B.measure[i];
        split[i] = (m == 0 || (m == 1 && Tj[Tp[i]] == i)) ? F_PT : OPEN;
    }
    for (int top = n - 1; top >= 0; --top) {
        const int c = B.node_at[top];
        --B.size[B.measure[c]];
        if (split[c] == F_PT) continue;
        split[c] = C_PT;
        for (int p = Tp[c]; p < Tp[c + 1]; ++p) {           // undecided nodes that depend on c become F ...
            const int f = Tj[p];
            if (split[f] != OPEN) continue;
            split[f] = F_PT;
            for (int q = Sp[f]; q < Sp[f + 1]; ++q) {       // ... and what they depend on gets more attractive
                const int k = Sj[q];
                if (split[k] == OPEN && B.measure[k] < n - 1) B.raise(k);
            }
        }
        for (int p = Sp[c]; p < Sp[c + 1]; ++p) {           // what c depends on loses one potential dependant
            const int j = Sj[p];
            if (split[j] == OPEN && B.measure[j] != 0) B.lower(j);
        }
    }
    return 0;
}

// amg_core.naive_aggregation (pyamg/amg_core/smoothed_aggregation.h:244-277; aggregate='naive'): in index order, a node
// not yet taken founds an aggregate with every neighbour not yet taken.  agg[i] in [0, n_agg), roots[k] = founder.
extern "C" int b200amg_naive_aggregation_host(int n, const int *Ap, const int *Aj, int *agg, int *roots)
{
    if (n < 0) return set_err(B200AMG_EINVAL, "naive_aggregation: n < 0");
    for (int i = 0; i < n; ++i) agg[i] = -1;
    int nagg = 0;
    for (int i = 0; i < n; ++i) {
        if (agg[i] >= 0) continue;
        agg[i] = nagg;
        for (int p = Ap[i]; p < Ap[i + 1]; ++p)
            if (agg[Aj[p]] < 0) agg[Aj[p]] = nagg;
        roots[nagg++] = i;
    }
    return nagg;
}

// ---- wavefronts of a lexicographic Gauss-Seidel sweep (amg_core.gauss_seidel / bsr_gauss_seidel / block_gauss_seidel,
// pyamg/amg_core/relaxation.h:33-62, :89-173, :753-806).  The sequential sweep over rows 0, 1, ..., n-1 relaxes row i
// with the NEW values of the rows j < i it references and the OLD values of the rows j > i.  level[i] = the earliest
// parallel step at which row i can run with exactly those operands: after every referenced row j < i (true dependence)
// and after every row j < i that references i (it must read x_i before i overwrites it).  Rows of one level are then
// independent of each other; running the levels in ascending order gives the bits of the forward sweep, in descending
// order those of the backward sweep.  One ascending pass; returns the number of levels.
extern "C" int b200amg_gs_wavefronts_host(int n, const int *Ap, const int *Aj, int *level)
{
    if (n < 0) return set_err(B200AMG_EINVAL, "gs_wavefronts: n < 0");
    std::vector<int> pending(n > 0 ? n : 1, 0);          // lower bound handed forward by rows that read x_i
    int nlev = 0;
    for (int i = 0; i < n; ++i) {
        int L = pending[i];
        for (int p = Ap[i]; p < Ap[i + 1]; ++p) {
            const int j = Aj[p];
            if (j >= 0 && j < i) L = std::max(L, level[j] + 1);
        }
        level[i] = L;
        for (int p = Ap[i]; p < Ap[i + 1]; ++p) {
            const int j = Aj[p];
            if (j > i && j < n) pending[j] = std::max(pending[j], L + 1);
        }
        nlev = std::max(nlev, L + 1);
    }
    return nlev;
}
