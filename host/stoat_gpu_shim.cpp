// stoat_gpu_shim.cpp — see stoat_gpu_shim.hpp.  Host plumbing only: flattening, ABI calls, TSV rows.
#include "stoat_gpu_shim.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>

namespace stoat_gpu_shim {

// ---- flattening ------------------------------------------------------------------------------------------------------
stoat_gpu_paths FlatPaths::view() const {
    stoat_gpu_paths p;
    p.n_snarls = snarl_path_off.size() - 1;
    p.snarl_path_off = snarl_path_off.data();
    p.path_edge_off = path_edge_off.data();
    p.edge_row = edge_row.data();
    return p;
}

FlatPaths flatten_snarl_paths(const std::vector<Snarl_data_t>& snarls, const EdgeBySampleMatrix& edge_matrix) {
    FlatPaths f;
    f.snarl_path_off.reserve(snarls.size() + 1);
    for (const Snarl_data_t& s : snarls) {
        for (const Path_traversal_t& path : s.snarl_paths) {
            const std::vector<Node_traversal_t>& nodes = path.get_paths();
            // consecutive node pairs are the path's edges (decompose_path_to_edges, src/snarl_analyzer.cpp:263-274)
            for (size_t i = 0; i + 1 < nodes.size(); ++i) {
                // an edge touching node 0 ('*', complex) is not checked (:328-330)
                if (nodes[i].get_node_id() == 0 || nodes[i + 1].get_node_id() == 0) continue;
                const size_t row = edge_matrix.find_edge(Edge_t(nodes[i], nodes[i + 1]));
                // an edge that is not in the matrix: the path matches nobody (:331-336)
                f.edge_row.push_back(row == std::numeric_limits<size_t>::max() ? STOAT_GPU_MISSING_EDGE : uint32_t(row));
            }
            f.path_edge_off.push_back(f.edge_row.size());
        }
        f.snarl_path_off.push_back(f.path_edge_off.size() - 1);
    }
    if (f.edge_row.empty()) f.edge_row.push_back(0);  // keep data() valid
    return f;
}

// ---- strings ---------------------------------------------------------------------------------------------------------
std::string set_precision(double value, uint8_t code) {
    char buf[64];
    if (stoat_gpu_format_value(value, code, buf, sizeof buf) < 0) throw std::runtime_error("stoat_gpu_format_value failed");
    return buf;
}

std::string format_group_paths(const uint32_t* g0, const uint32_t* g1, size_t n) {
    std::string out;
    for (size_t j = 0; j < n; ++j) {
        if (j) out += ',';
        out += std::to_string(g0[j]) + ":" + std::to_string(g1[j]);
    }
    return out;
}

std::string type_variants_string(const Snarl_data_t& s) {
    std::string out;
    for (size_t i = 0; i < s.type_variants.size(); ++i) {
        if (i) out += ',';
        out += s.type_variants[i];
    }
    return out;
}

std::vector<size_t> found_gene_snarl(const std::vector<Qtl_data_t>& genes, size_t start, size_t end, size_t window) {
    std::vector<size_t> idx;
    const size_t lo = start > window ? start - window : 0;
    const size_t hi = end + window;
    for (size_t i = 0; i < genes.size(); ++i)
        if (!(genes[i].end_pos < lo || genes[i].start_pos > hi)) idx.push_back(i);  // overlap test, :485
    return idx;
}

namespace {
std::string snarl_name(const Snarl_data_t& s) {  // stoat::pairToString, src/snarl_data_t.cpp:181-185
    return std::to_string(s.snarl_ids.first) + "_" + std::to_string(s.snarl_ids.second);
}
std::string join(const uint32_t* v, size_t n) {  // stoat::vectorToString, src/utils.cpp:102-110
    std::string out;
    for (size_t i = 0; i < n; ++i) {
        if (i) out += ',';
        out += std::to_string(v[i]);
    }
    return out;
}
void row_prefix(std::ostream& o, const std::string& chr, const Snarl_data_t& s) {
    o << chr << "\t" << s.start_positions << "\t" << s.end_positions << "\t" << snarl_name(s) << "\t"
      << type_variants_string(s) << "\t";
}
std::vector<double> flatten_covariates(const std::vector<std::vector<double>>& covariate, size_t& n_covar) {
    n_covar = covariate.empty() ? 0 : covariate[0].size();
    std::vector<double> flat;
    flat.reserve(covariate.size() * n_covar);
    for (const auto& r : covariate) {
        if (r.size() != n_covar) throw std::runtime_error("covariate rows of unequal length");
        flat.insert(flat.end(), r.begin(), r.end());
    }
    return flat;
}
}  // namespace

// ---- base ------------------------------------------------------------------------------------------------------------
GpuSnarlAnalyzer::GpuSnarlAnalyzer(const EdgeBySampleMatrix& m, size_t min_individuals, size_t min_haplotypes,
                                   double maf_threshold, int device)
    : edge_matrix(m) {
    thr.min_individuals = min_individuals;
    thr.min_haplotypes = min_haplotypes;
    thr.maf_threshold = maf_threshold;
    const int rc = stoat_gpu_create(&ctx, device);
    if (rc != STOAT_GPU_OK) throw std::runtime_error(std::string("stoat_gpu_create: ") + stoat_gpu_last_error(nullptr));
}

GpuSnarlAnalyzer::~GpuSnarlAnalyzer() { stoat_gpu_destroy(ctx); }

void GpuSnarlAnalyzer::check(int rc, const char* what) const {
    if (rc != STOAT_GPU_OK) throw std::runtime_error(std::string(what) + ": " + stoat_gpu_last_error(ctx));
}

stoat_gpu_timing GpuSnarlAnalyzer::last_timing() const {
    stoat_gpu_timing t{};
    stoat_gpu_last_timing(ctx, &t);
    return t;
}

size_t GpuSnarlAnalyzer::analyze_and_write_chunk(const std::vector<Snarl_data_t>& snarls, const std::string& chr,
                                                 std::ostream& outf) {
    if (snarls.empty()) return 0;
    check(stoat_gpu_set_matrix(ctx, edge_matrix.data().data(), edge_matrix.rows(), edge_matrix.cols()), "set_matrix");
    upload_phenotype();
    const FlatPaths flat = flatten_snarl_paths(snarls, edge_matrix);
    return run(snarls, flat, chr, outf);
}

// ---- binary ----------------------------------------------------------------------------------------------------------
GpuBinarySnarlAnalyzer::GpuBinarySnarlAnalyzer(const EdgeBySampleMatrix& m, const std::vector<bool>& pheno,
                                               size_t min_individuals, size_t min_haplotypes, double maf_threshold,
                                               int device)
    : GpuSnarlAnalyzer(m, min_individuals, min_haplotypes, maf_threshold, device), binary_phenotype(pheno) {}

void GpuBinarySnarlAnalyzer::write_header(std::ostream& outf) const {
    outf << "#CHR\tSTART_POS\tEND_POS\tSNARL\tPATH_LENGTHS\tP_FISHER\tP_CHI2\tGROUP_PATHS\tDEPTH\n";
}

void GpuBinarySnarlAnalyzer::upload_phenotype() {
    std::vector<uint8_t> is_case(binary_phenotype.size());
    for (size_t i = 0; i < is_case.size(); ++i) is_case[i] = binary_phenotype[i] ? 1 : 0;
    check(stoat_gpu_set_binary_phenotype(ctx, is_case.data(), is_case.size()), "set_binary_phenotype");
}

size_t GpuBinarySnarlAnalyzer::run(const std::vector<Snarl_data_t>& snarls, const FlatPaths& flat, const std::string& chr,
                                   std::ostream& outf) {
    const size_t S = snarls.size(), P = flat.path_edge_off.size() - 1;
    std::vector<uint32_t> g0(P), g1(P);
    std::vector<uint64_t> total(S);
    std::vector<uint8_t> filtered(S), chi2_code(S), fisher_code(S);
    std::vector<double> p_chi2(S), p_fisher(S);
    stoat_gpu_binary_out out{g0.data(), g1.data(), total.data(), filtered.data(), p_chi2.data(), p_fisher.data(),
                             chi2_code.data(), fisher_code.data()};
    const stoat_gpu_paths pv = flat.view();
    check(stoat_gpu_binary_batch(ctx, &pv, &thr, &out), "binary_batch");
    size_t n_filtered = 0;
    std::vector<uint32_t> c0, c1;
    for (size_t s = 0; s < S; ++s) {
        if (filtered[s]) {
            ++n_filtered;
            continue;
        }
        // remove_empty_columns_binary_table (src/snarl_analyzer.cpp:686-702): GROUP_PATHS prints the kept columns
        c0.clear();
        c1.clear();
        for (uint64_t p = flat.snarl_path_off[s]; p < flat.snarl_path_off[s + 1]; ++p)
            if (g0[p] + g1[p] != 0) {
                c0.push_back(g0[p]);
                c1.push_back(g1[p]);
            }
        row_prefix(outf, chr, snarls[s]);  // stoat::write_binary, src/writer.cpp:23-35
        outf << set_precision(p_fisher[s], fisher_code[s]) << "\t" << set_precision(p_chi2[s], chi2_code[s]) << "\t"
             << format_group_paths(c0.data(), c1.data(), c0.size()) << "\t" << snarls[s].depth << "\n";
    }
    return n_filtered;
}

// ---- quantitative ----------------------------------------------------------------------------------------------------
GpuQuantitativeSnarlAnalyzer::GpuQuantitativeSnarlAnalyzer(const EdgeBySampleMatrix& m, const std::vector<double>& pheno,
                                                           const std::vector<std::vector<double>>& covar,
                                                           size_t min_individuals, size_t min_haplotypes,
                                                           double maf_threshold, int device)
    : GpuSnarlAnalyzer(m, min_individuals, min_haplotypes, maf_threshold, device), quantitative_phenotype(pheno),
      covariate(covar) {}

void GpuQuantitativeSnarlAnalyzer::write_header(std::ostream& outf) const {
    outf << "#CHR\tSTART_POS\tEND_POS\tSNARL\tPATH_LENGTHS\tP\tRSQUARE\tBETA\tSE\tALLELE_PATHS\tDEPTH\n";
}

void GpuQuantitativeSnarlAnalyzer::upload_phenotype() {
    check(stoat_gpu_set_quantitative_phenotype(ctx, quantitative_phenotype.data(), quantitative_phenotype.size()),
          "set_quantitative_phenotype");
    size_t n_covar = 0;
    const std::vector<double> flat = flatten_covariates(covariate, n_covar);
    check(stoat_gpu_set_covariates(ctx, n_covar ? flat.data() : nullptr, covariate.size(), n_covar), "set_covariates");
}

namespace {
void check_ols_code(uint8_t code, const Snarl_data_t& s) {
    if (code == STOAT_P_OVERLIMIT)
        throw std::runtime_error("snarl " + snarl_name(s) + " exceeds STOAT_GPU_MAX_PATHS_PER_SNARL / "
                                 "STOAT_GPU_MAX_OLS_FEATURES: run the per-snarl host analyzer for it");
}
}  // namespace

size_t GpuQuantitativeSnarlAnalyzer::run(const std::vector<Snarl_data_t>& snarls, const FlatPaths& flat,
                                         const std::string& chr, std::ostream& outf) {
    const size_t S = snarls.size(), P = flat.path_edge_off.size() - 1;
    std::vector<uint32_t> allele(P), n_used(S);
    std::vector<uint8_t> filtered(S), code(S);
    std::vector<double> p(S), r2(S), beta(S), se(S);
    stoat_gpu_quant_out out{allele.data(), filtered.data(), n_used.data(), p.data(), r2.data(), beta.data(), se.data(),
                            code.data()};
    const stoat_gpu_paths pv = flat.view();
    check(stoat_gpu_quantitative_batch(ctx, &pv, &thr, &out), "quantitative_batch");
    size_t n_filtered = 0;
    for (size_t s = 0; s < S; ++s) {
        if (filtered[s]) {
            ++n_filtered;
            continue;
        }
        check_ols_code(code[s], snarls[s]);
        const uint64_t p0 = flat.snarl_path_off[s], np = flat.snarl_path_off[s + 1] - p0;
        row_prefix(outf, chr, snarls[s]);  // stoat::write_quantitative, src/writer.cpp:53-68
        outf << set_precision(p[s]) << "\t" << set_precision(r2[s]) << "\t" << set_precision(beta[s]) << "\t"
             << set_precision(se[s]) << "\t" << join(allele.data() + p0, np) << "\t" << snarls[s].depth << "\n";
    }
    return n_filtered;
}

// ---- eQTL ------------------------------------------------------------------------------------------------------------
GpuEQTLSnarlAnalyzer::GpuEQTLSnarlAnalyzer(const EdgeBySampleMatrix& m,
                                           const std::unordered_map<std::string, std::vector<Qtl_data_t>>& map,
                                           const std::vector<std::vector<double>>& covar, size_t window,
                                           size_t min_individuals, size_t min_haplotypes, double maf_threshold, int device)
    : GpuSnarlAnalyzer(m, min_individuals, min_haplotypes, maf_threshold, device), eqtl_map(map), covariate(covar),
      windows_gene_threshold(window) {}

void GpuEQTLSnarlAnalyzer::write_header(std::ostream& outf) const {
    outf << "#CHR\tSTART_POS\tEND_POS\tSNARL\tPATH_LENGTHS\tGENE\tP\tRSQUARE\tBETA\tSE\tALLELE_PATHS\tDEPTH\n";
}

void GpuEQTLSnarlAnalyzer::upload_phenotype() {
    size_t n_covar = 0;
    const std::vector<double> flat = flatten_covariates(covariate, n_covar);
    check(stoat_gpu_set_covariates(ctx, n_covar ? flat.data() : nullptr, covariate.size(), n_covar), "set_covariates");
}

size_t GpuEQTLSnarlAnalyzer::run(const std::vector<Snarl_data_t>& snarls, const FlatPaths& flat, const std::string& chr,
                                 std::ostream& outf) {
    const std::vector<Qtl_data_t>& genes = eqtl_map.at(chr);  // throws like the reference's .at() (:496)
    const size_t S = snarls.size(), P = flat.path_edge_off.size() - 1, G = genes.size();
    const size_t n = edge_matrix.cols() / 2;
    // gene windows on the host (found_gene_snarl, :471-491) -> CSR
    std::vector<uint64_t> off{0};
    std::vector<uint32_t> idx;
    for (const Snarl_data_t& s : snarls) {
        for (size_t g : found_gene_snarl(genes, s.start_positions, s.end_positions, windows_gene_threshold))
            idx.push_back(uint32_t(g));
        off.push_back(idx.size());
    }
    std::vector<double> expr(G * n);
    for (size_t g = 0; g < G; ++g) {
        if (genes[g].sampleExpresion.size() != n) throw std::runtime_error("expression vector length != samples");
        std::copy(genes[g].sampleExpresion.begin(), genes[g].sampleExpresion.end(), expr.begin() + g * n);
    }
    const size_t n_pairs = idx.size();
    if (idx.empty()) idx.push_back(0);
    std::vector<uint32_t> allele(P), n_used(S);
    std::vector<uint8_t> filtered(S), code(std::max<size_t>(n_pairs, 1));
    std::vector<double> p(code.size()), r2(code.size()), beta(code.size()), se(code.size());
    stoat_gpu_eqtl_genes gv{G, expr.data(), off.data(), idx.data()};
    stoat_gpu_eqtl_out out{allele.data(), filtered.data(), n_used.data(), p.data(), r2.data(), beta.data(), se.data(),
                           code.data()};
    const stoat_gpu_paths pv = flat.view();
    check(stoat_gpu_eqtl_batch(ctx, &pv, &gv, &thr, &out), "eqtl_batch");
    size_t n_filtered = 0;
    for (size_t s = 0; s < S; ++s) {
        if (filtered[s]) {
            ++n_filtered;
            continue;
        }
        const uint64_t p0 = flat.snarl_path_off[s], np = flat.snarl_path_off[s + 1] - p0;
        for (uint64_t k = off[s]; k < off[s + 1]; ++k) {
            check_ols_code(code[k], snarls[s]);
            row_prefix(outf, chr, snarls[s]);  // stoat::write_eqtl, src/writer.cpp:70-87
            outf << genes[idx[k]].geneName << "\t" << set_precision(p[k]) << "\t" << set_precision(r2[k]) << "\t"
                 << set_precision(beta[k]) << "\t" << set_precision(se[k]) << "\t" << join(allele.data() + p0, np) << "\t"
                 << snarls[s].depth << "\n";
        }
    }
    return n_filtered;
}

}  // namespace stoat_gpu_shim
