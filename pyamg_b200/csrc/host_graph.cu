// Host-side setup helpers that are inherently sequential in the reference and must give the
// IDENTICAL integer result: greedy MIS colouring (pyamg/amg_core/graph.h:171-188 over
// :34-61) used to build the multicolour Gauss-Seidel ordering, and the three-pass standard
// aggregation (pyamg/amg_core/smoothed_aggregation.h:121-221).  Arrays are HOST pointers.
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
