// host_select.cu — host-side top-gene selection (no device work): the `top` genes by
// (max isoform coverage, gene id), every isoform of a selected gene kept.  Exact restatement of
// the reference's TopGeneSelector.get_top_genes (select_top_genes.py:108-127: defaultdict(float)
// maxima, heapq.nlargest over (coverage, gene) tuples, so ties resolve to the larger gene id).
#include <algorithm>
#include <unordered_map>
#include <utility>
#include <vector>
#include "common.cuh"

extern "C" int64_t rcq_top_gene_mask(const double* coverage, const int32_t* gene, int64_t n, int64_t top,
                                     uint8_t* keep) {
    if (n < 0 || (n > 0 && (!coverage || !gene || !keep))) return rcq_fail(RCQ_ERR_ARG, "rcq_top_gene_mask: bad argument");
    if (n == 0) return 0;
    int32_t gmin = gene[0], gmax = gene[0];
    for (int64_t i = 1; i < n; i++) { gmin = std::min(gmin, gene[i]); gmax = std::max(gmax, gene[i]); }
    std::vector<std::pair<double, int32_t>> genes;   // (best coverage, gene id), one per distinct gene
    genes.reserve((size_t)n);
    std::vector<int32_t> slot(n);                    // transcript -> index into `genes`
    const int64_t span = (int64_t)gmax - gmin + 1;
    if (span <= 8 * n + 1024) {
        std::vector<int32_t> at((size_t)span, -1);
        for (int64_t i = 0; i < n; i++) {
            int32_t& a = at[(size_t)((int64_t)gene[i] - gmin)];
            if (a < 0) { a = (int32_t)genes.size(); genes.emplace_back(0.0, gene[i]); }
            slot[i] = a;
            if (coverage[i] > genes[a].first) genes[a].first = coverage[i];
        }
    } else {
        std::unordered_map<int32_t, int32_t> at;
        at.reserve((size_t)n * 2);
        for (int64_t i = 0; i < n; i++) {
            auto it = at.find(gene[i]);
            int32_t a;
            if (it == at.end()) { a = (int32_t)genes.size(); at.emplace(gene[i], a); genes.emplace_back(0.0, gene[i]); }
            else a = it->second;
            slot[i] = a;
            if (coverage[i] > genes[a].first) genes[a].first = coverage[i];
        }
    }
    const int64_t G = (int64_t)genes.size();
    std::vector<uint8_t> sel((size_t)G, 0);
    if (top >= G) std::fill(sel.begin(), sel.end(), 1);
    else if (top > 0) {
        // the top-th largest (coverage, gene id) pair by selection on a contiguous copy (no index indirection); gene
        // ids are distinct, so exactly `top` genes compare >= that pair
        std::vector<std::pair<double, int32_t>> ord(genes);
        std::nth_element(ord.begin(), ord.begin() + (G - top), ord.end());
        const std::pair<double, int32_t> thr = ord[(size_t)(G - top)];
        for (int64_t g = 0; g < G; g++) sel[g] = !(genes[g] < thr);
    }
    int64_t kept = 0;
    for (int64_t i = 0; i < n; i++) { keep[i] = sel[slot[i]]; kept += keep[i]; }
    return kept;
}
