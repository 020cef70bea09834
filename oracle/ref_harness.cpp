// ref_harness.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A thin driver around the REFERENCE'S OWN hot-path code.  oracle/ref_build.py compiles this file together with the
// reference sources where they lie under /root/reference/src (unchanged, never copied) into oracle/_ref/libstoat_ref.so.
// What runs below is therefore the reference's BinarySnarlAnalyzer / QuantitativeSnarlAnalyzer / EQTLSnarlAnalyzer
// ::analyze_and_write_snarl (src/snarl_analyzer.cpp:358-390, 431-467, 493-539), its tables, filters, FisherKhi2,
// LinearRegression (vendored Eigen) and its TSV writer; only Boost's distribution tails come from the stand-ins in
// oracle/ref_stubs/ (Boost is a system package absent from this image).
//
// The harness does what `stoat vcf` does around that code and nothing more:
//   * fills an EdgeBySampleMatrix (bytes in the reference layout + the edge -> row header) in place of
//     make_edge_matrix (src/snarl_analyzer.cpp:190-260, needs htslib + a VCF);
//   * builds the Snarl_data_t records in place of parse_snarl_path (src/snarl_data_t.cpp:40-110);
//   * runs the OpenMP loop of process_snarls_by_chromosome_chunk (src/snarl_analyzer.cpp:161-172) and times it the way
//     the reference's "Snarl time analysis" log line does (:162, 175-178).
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include <omp.h>

#include "snarl_analyzer.hpp"
#include "stats_test.hpp"
#include "utils.hpp"
#include "writer.hpp"

// Stand-in for src/io/register_io.cpp (needs libvgio/libbdsg; graph loading is outside the path, never called here).
namespace stoat { namespace io {
bool register_libvg_io() { throw std::logic_error("oracle/_ref: libvgio stand-in called"); }
}}  // namespace stoat::io

namespace {

// EdgeBySampleMatrix keeps its storage protected (src/matrix.hpp:51-55); a derived class may fill it.
struct FilledMatrix : stoat_vcf::EdgeBySampleMatrix {
    FilledMatrix(const std::vector<std::string>& names, size_t rows, size_t cols)
        : stoat_vcf::EdgeBySampleMatrix(names, rows, cols) {}
    void fill(const uint8_t* bits, uint64_t n_rows, uint64_t n_cols, const uint64_t* edge_desc) {
        const size_t n_bytes = (n_rows * n_cols + 7) / 8;
        matrix_1D.assign(bits, bits + n_bytes);
        row_header.clear();
        row_header.reserve(n_rows);
        for (uint64_t r = 0; r < n_rows; ++r) {
            const uint64_t* e = edge_desc + 4 * r;
            row_header.emplace(stoat::Edge_t(stoat::Node_traversal_t(e[0], e[1] != 0), stoat::Node_traversal_t(e[2], e[3] != 0)), r);
        }
    }
};

struct Chunk {
    std::vector<std::string> samples;
    std::unordered_map<std::string, std::vector<stoat::Snarl_data_t>> chr_to_snarls;
};

const char* kChr = "chr1";

void make_snarls(Chunk& c, uint64_t n_samples, uint64_t n_snarls, const uint64_t* snarl_path_off, const uint64_t* path_node_off,
                 const uint64_t* node_id, const uint8_t* node_rev, const uint64_t* snarl_pos) {
    c.samples.resize(n_samples);
    for (uint64_t i = 0; i < n_samples; ++i) c.samples[i] = "s" + std::to_string(i);
    auto& v = c.chr_to_snarls[kChr];
    v.reserve(n_snarls);
    for (uint64_t s = 0; s < n_snarls; ++s) {
        std::vector<stoat::Path_traversal_t> paths;
        std::vector<std::string> types;
        for (uint64_t p = snarl_path_off[s]; p < snarl_path_off[s + 1]; ++p) {
            stoat::Path_traversal_t pt;
            for (uint64_t k = path_node_off[p]; k < path_node_off[p + 1]; ++k)
                pt.add_node_traversal_t(stoat::Node_traversal_t(node_id[k], node_rev[k] != 0));
            paths.push_back(pt);
            types.push_back("1");
        }
        const uint64_t start = snarl_pos ? snarl_pos[2 * s] : s, end = snarl_pos ? snarl_pos[2 * s + 1] : s;
        // SNARL column = "<s>_<s>": the tests key rows by it (row order under OpenMP is not deterministic)
        v.emplace_back(handlegraph::net_handle_t(), std::make_pair(size_t(s), size_t(s)), paths, size_t(start), size_t(end), types, size_t(1));
    }
}

// the loop of SnarlAnalyzer::process_snarls_by_chromosome_chunk (src/snarl_analyzer.cpp:161-172)
double run_loop(stoat_vcf::SnarlAnalyzer& an, const std::vector<stoat::Snarl_data_t>& snarls, int n_threads, const char* tsv,
                uint64_t* n_filtered) {
    std::ofstream outf(tsv);
    an.write_header(outf);
    const std::string chr = kChr;
    uint64_t filtered = 0;
    if (n_threads > 0) omp_set_num_threads(n_threads);
    const auto t0 = std::chrono::high_resolution_clock::now();
#pragma omp parallel for schedule(static) reduction(+ : filtered)
    for (size_t i = 0; i < snarls.size(); ++i) {
        const bool fail = an.analyze_and_write_snarl(snarls[i], chr, outf);
        filtered += fail ? 1 : 0;
    }
    const auto t1 = std::chrono::high_resolution_clock::now();
    outf.close();
    if (n_filtered) *n_filtered = filtered;
    return std::chrono::duration<double>(t1 - t0).count();
}

std::vector<std::vector<double>> covar_rows(const double* covar, uint64_t n, uint64_t nc) {
    std::vector<std::vector<double>> out;
    if (!covar || nc == 0) return out;
    out.resize(n);
    for (uint64_t i = 0; i < n; ++i) out[i].assign(covar + i * nc, covar + (i + 1) * nc);
    return out;
}

int put(const std::string& s, char* buf, uint64_t len) {
    if (s.size() + 1 > len) return -1;
    memcpy(buf, s.c_str(), s.size() + 1);
    return int(s.size());
}

}  // namespace

extern "C" {

// trait: 0 binary (chi2 + Fisher), 1 quantitative (OLS), 2 eQTL (OLS per gene of the window), 3 binary with covariates
// (logistic regression)
// edge_desc [n_rows][4] = {id1, rev1, id2, rev2}; snarl_pos [n_snarls][2] or NULL; thr = {min_individuals, min_haplotypes}
double ref_run(int trait, const uint8_t* bits, uint64_t n_rows, uint64_t n_cols, const uint64_t* edge_desc, uint64_t n_snarls,
               const uint64_t* snarl_path_off, const uint64_t* path_node_off, const uint64_t* node_id, const uint8_t* node_rev,
               const uint64_t* snarl_pos, const uint8_t* is_case, const double* y, const double* covar, uint64_t n_covar,
               const double* expression, uint64_t n_genes, const uint64_t* gene_pos, uint64_t window, uint64_t min_individuals,
               uint64_t min_haplotypes, double maf, int n_threads, const char* tsv_path, uint64_t* n_filtered) {
    const uint64_t n_samples = n_cols / 2;
    Chunk c;
    make_snarls(c, n_samples, n_snarls, snarl_path_off, path_node_off, node_id, node_rev, snarl_pos);
    FilledMatrix m(c.samples, n_rows, n_cols);
    m.fill(bits, n_rows, n_cols, edge_desc);
    const double table_threshold = -1;
    const size_t min_ind = min_individuals, min_hap = min_haplotypes;
    const std::string regression_dir;
    const auto cov = covar_rows(covar, n_samples, n_covar);
    const auto& snarls = c.chr_to_snarls.at(kChr);
    if (trait == 0) {
        std::vector<bool> pheno(n_samples);
        for (uint64_t i = 0; i < n_samples; ++i) pheno[i] = is_case[i] != 0;
        stoat_vcf::BinarySnarlAnalyzer an(c.chr_to_snarls, m, c.samples, maf, table_threshold, pheno, min_ind, min_hap, regression_dir);
        return run_loop(an, snarls, n_threads, tsv_path, n_filtered);
    }
    if (trait == 1) {
        std::vector<double> pheno(y, y + n_samples);
        stoat_vcf::QuantitativeSnarlAnalyzer an(c.chr_to_snarls, m, c.samples, cov, maf, table_threshold, pheno, min_ind, min_hap,
                                                regression_dir);
        return run_loop(an, snarls, n_threads, tsv_path, n_filtered);
    }
    if (trait == 3) {  // BinaryCovarSnarlAnalyzer (src/snarl_analyzer.cpp:392-428): logistic regression
        std::vector<bool> pheno(n_samples);
        for (uint64_t i = 0; i < n_samples; ++i) pheno[i] = is_case[i] != 0;
        stoat_vcf::BinaryCovarSnarlAnalyzer an(c.chr_to_snarls, m, c.samples, cov, maf, table_threshold, pheno, min_ind, min_hap,
                                               regression_dir);
        return run_loop(an, snarls, n_threads, tsv_path, n_filtered);
    }
    if (trait == 2) {
        std::unordered_map<std::string, std::vector<stoat_vcf::Qtl_data>> eqtl;
        auto& genes = eqtl[kChr];
        genes.reserve(n_genes);
        for (uint64_t g = 0; g < n_genes; ++g) {
            std::vector<double> e(expression + g * n_samples, expression + (g + 1) * n_samples);
            genes.emplace_back("g" + std::to_string(g), e, size_t(gene_pos ? gene_pos[2 * g] : 0), size_t(gene_pos ? gene_pos[2 * g + 1] : 0));
        }
        const size_t win = window;
        stoat_vcf::EQTLSnarlAnalyzer an(c.chr_to_snarls, m, c.samples, cov, maf, table_threshold, eqtl, win, min_ind, min_hap,
                                        regression_dir);
        return run_loop(an, snarls, n_threads, tsv_path, n_filtered);
    }
    return -1.0;
}

// LinearRegression::linear_regression (src/stats_test.cpp:424-506) -> "p\tbeta\tse\tr2"
int ref_linear_regression(const double* df_rowmajor, uint64_t n, uint64_t nv, const double* y, const double* covar, uint64_t nc,
                          char* buf, uint64_t len) {
    std::vector<std::vector<double>> df(n);
    for (uint64_t i = 0; i < n; ++i) df[i].assign(df_rowmajor + i * nv, df_rowmajor + (i + 1) * nv);
    std::vector<double> yy(y, y + n);
    stoat::LinearRegression lr;
    const auto [p, beta, se, r2] = lr.linear_regression(df, yy, covar_rows(covar, n, nc));
    return put(p + "\t" + beta + "\t" + se + "\t" + r2, buf, len);
}

// LogisticRegression::logistic_regression (src/stats_test.cpp:49-176) -> "p\tbeta\tse"
int ref_logistic_regression(const double* df_rowmajor, uint64_t n, uint64_t nv, const uint8_t* y, const double* covar, uint64_t nc,
                            char* buf, uint64_t len) {
    std::vector<std::vector<double>> df(n);
    for (uint64_t i = 0; i < n; ++i) df[i].assign(df_rowmajor + i * nv, df_rowmajor + (i + 1) * nv);
    std::vector<bool> yy(n);
    for (uint64_t i = 0; i < n; ++i) yy[i] = y[i] != 0;
    stoat::LogisticRegression lr;
    const auto [p, beta, se] = lr.logistic_regression(df, yy, covar_rows(covar, n, nc));
    return put(p + "\t" + beta + "\t" + se, buf, len);
}

// FisherKhi2::fisher_khi2 (src/stats_test.cpp:362-380) -> "chi2\tfisher"
int ref_fisher_khi2(const uint64_t* g0, const uint64_t* g1, uint64_t n, char* buf, uint64_t len) {
    std::vector<size_t> a(g0, g0 + n), b(g1, g1 + n);
    stoat::FisherKhi2 fk;
    const auto [chi2, fisher] = fk.fisher_khi2(a, b);
    return put(chi2 + "\t" + fisher, buf, len);
}

int ref_chi2_2x2(uint64_t a, uint64_t b, uint64_t c, uint64_t d, char* buf, uint64_t len) {
    stoat::FisherKhi2 fk;
    return put(fk.chi2_2x2(a, b, c, d), buf, len);
}

int ref_fisher_2x2(uint64_t a, uint64_t b, uint64_t c, uint64_t d, char* buf, uint64_t len) {
    stoat::FisherKhi2 fk;
    return put(fk.fastFishersExactTest(a, b, c, d), buf, len);
}

int ref_set_precision(double v, char* buf, uint64_t len) { return put(stoat::set_precision(v), buf, len); }

int ref_num_threads(void) { return omp_get_max_threads(); }

}  // extern "C"
