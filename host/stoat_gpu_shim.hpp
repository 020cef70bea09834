// stoat_gpu_shim.hpp — C++ host side of the drop-in boundary.
//
// The reference has no plugin API: its hot loop is the body of
//   SnarlAnalyzer::process_snarls_by_chromosome_chunk            src/snarl_analyzer.cpp:161-172
// which calls the virtual analyze_and_write_snarl (src/snarl_analyzer.hpp:66) once per snarl.  The classes below are
// the batched stand-ins of BinarySnarlAnalyzer (:358-390), QuantitativeSnarlAnalyzer (:431-467) and EQTLSnarlAnalyzer
// (:493-539): same constructor arguments, same thresholds, same TSV rows, but ONE call per chromosome chunk
// (analyze_and_write_chunk) that flattens the snarls' paths, calls the C ABI (include/stoat_gpu.h) and writes the
// rows in snarl order.  Errors of the ABI become std::runtime_error, as every other failure in the reference is an
// exception that reaches main().  There is no CPU path here: without libstoat_gpu.so and a B200 the calls throw.
//
// Inside the reference tree compile with -DSTOAT_GPU_SHIM_USE_REFERENCE_TYPES: the shim then uses stoat::Snarl_data_t,
// stoat::Edge_t and stoat_vcf::EdgeBySampleMatrix themselves (the latter needs the two accessors INTEGRATION.md adds).
// Stand-alone (this repository's tests) it uses the look-alike types below, which expose exactly the members the
// shim touches, with the reference's names and meaning.
#ifndef STOAT_GPU_SHIM_HPP
#define STOAT_GPU_SHIM_HPP

#include <cstddef>
#include <cstdint>
#include <limits>
#include <ostream>
#include <set>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../include/stoat_gpu.h"

#ifdef STOAT_GPU_SHIM_USE_REFERENCE_TYPES
#include "matrix.hpp"
#include "snarl_data_t.hpp"
namespace stoat_gpu_shim {
using stoat::Edge_t;
using stoat::Node_traversal_t;
using stoat::Path_traversal_t;
using stoat::Snarl_data_t;
using stoat_vcf::EdgeBySampleMatrix;
}  // namespace stoat_gpu_shim
#else
namespace stoat_gpu_shim {

// node with orientation — src/snarl_data_t.hpp:40-58
struct Node_traversal_t {
    Node_traversal_t(const size_t& id, const bool& rev) : node_id(id), is_reverse(rev) {}
    size_t get_node_id() const { return node_id; }
    bool get_is_reverse() const { return is_reverse; }
    bool operator==(const Node_traversal_t& o) const { return node_id == o.node_id && is_reverse == o.is_reverse; }

private:
    size_t node_id : 63;
    bool is_reverse : 1;
};

// edge between two oriented nodes — src/snarl_data_t.hpp:61-78
struct Edge_t {
    Edge_t(const Node_traversal_t& a, const Node_traversal_t& b) : edge(a, b) {}
    std::pair<size_t, size_t> print_pair_edge() const { return {edge.first.get_node_id(), edge.second.get_node_id()}; }
    const std::pair<Node_traversal_t, Node_traversal_t>& get_edge() const { return edge; }
    bool operator==(const Edge_t& o) const { return edge == o.edge; }

private:
    std::pair<Node_traversal_t, Node_traversal_t> edge;
};

// path through a snarl — src/snarl_data_t.hpp:81-96
struct Path_traversal_t {
    void add_node_traversal_t(const Node_traversal_t& n) { paths.push_back(n); }
    const std::vector<Node_traversal_t>& get_paths() const { return paths; }

private:
    std::vector<Node_traversal_t> paths;
};

// the fields of stoat::Snarl_data_t (src/snarl_data_t.hpp:99-118) that the analyzers and the writer read
struct Snarl_data_t {
    std::vector<std::string> type_variants;
    std::vector<Path_traversal_t> snarl_paths;
    std::pair<size_t, size_t> snarl_ids;
    size_t start_positions = 0;
    size_t end_positions = 0;
    size_t depth = 0;
};

struct EdgeHash {
    size_t operator()(const Edge_t& e) const {
        const auto& p = e.get_edge();
        uint64_t a = (uint64_t(p.first.get_node_id()) << 1) | uint64_t(p.first.get_is_reverse());
        uint64_t b = (uint64_t(p.second.get_node_id()) << 1) | uint64_t(p.second.get_is_reverse());
        uint64_t x = a * 0x9E3779B97F4A7C15ull ^ (b + 0xBF58476D1CE4E5B9ull);
        x ^= x >> 31;
        return size_t(x * 0x94D049BB133111EBull);
    }
};

// Bit matrix, rows = edges, columns = haplotypes — the reference's layout (src/matrix.cpp:67-84): bit index
// row * cols + col, least significant bit first, rows NOT byte aligned; rows are created on first use
// (push_matrix, :40-64) and the storage doubles when full (:26-31).
class EdgeBySampleMatrix {
public:
    EdgeBySampleMatrix(const std::vector<std::string>& names, size_t rows, size_t cols)
        : cols_(cols), MaxElement(rows * cols), matrix_1D((rows * cols + 7) / 8, 0), sampleNames(names) {}
    bool operator()(size_t row, size_t col) const {
        const size_t bit = row * cols_ + col;
        return (matrix_1D[bit / 8] >> (bit % 8)) & 1u;
    }
    void set(size_t row, size_t col) {
        const size_t bit = row * cols_ + col;
        matrix_1D[bit / 8] |= uint8_t(1u << (bit % 8));
    }
    size_t getMaxElement() const { return MaxElement; }
    void expandMatrix() {
        MaxElement *= 2;
        matrix_1D.resize((MaxElement + 7) / 8, 0);
    }
    size_t find_edge(const Edge_t& edge) const {
        const auto it = row_header.find(edge);
        return it == row_header.end() ? std::numeric_limits<size_t>::max() : it->second;
    }
    void push_matrix(const Edge_t& edge, size_t col) {
        const auto it = row_header.emplace(edge, row_header.size()).first;
        while ((it->second + 1) * cols_ > MaxElement) expandMatrix();
        set(it->second, col);
    }
    // the two accessors INTEGRATION.md adds to the reference class
    const std::vector<uint8_t>& data() const { return matrix_1D; }
    size_t cols() const { return cols_; }
    size_t rows() const { return row_header.size(); }

protected:
    size_t cols_;
    size_t MaxElement;
    std::vector<uint8_t> matrix_1D;
    std::unordered_map<Edge_t, size_t, EdgeHash> row_header;

public:
    std::vector<std::string> sampleNames;
};

}  // namespace stoat_gpu_shim
#endif  // STOAT_GPU_SHIM_USE_REFERENCE_TYPES

namespace stoat_gpu_shim {

// Qtl_data_t of the reference (src/arg_parser.hpp): one gene of a chromosome with its expression per sample
struct Qtl_data_t {
    std::string geneName;
    std::vector<double> sampleExpresion;  // (sic) the reference's spelling
    size_t start_pos = 0;
    size_t end_pos = 0;
};

// CSR of the checked matrix rows of every path of a chunk of snarls: decompose_path_to_edges
// (src/snarl_analyzer.cpp:263-274) + the front half of identify_path (:321-337)
struct FlatPaths {
    std::vector<uint64_t> snarl_path_off{0};
    std::vector<uint64_t> path_edge_off{0};
    std::vector<uint32_t> edge_row;
    stoat_gpu_paths view() const;
};
FlatPaths flatten_snarl_paths(const Snarl_data_t* snarls, size_t n_snarls, const EdgeBySampleMatrix& edge_matrix);
inline FlatPaths flatten_snarl_paths(const std::vector<Snarl_data_t>& snarls, const EdgeBySampleMatrix& edge_matrix) {
    return flatten_snarl_paths(snarls.data(), snarls.size(), edge_matrix);
}

// the reference's output strings
std::string set_precision(double value, uint8_t code = STOAT_P_VALUE);                          // src/utils.cpp:5-15
std::string format_group_paths(const uint32_t* g0, const uint32_t* g1, size_t n);                // src/binary_table.cpp:6-17
std::string type_variants_string(const Snarl_data_t& s);                                         // src/snarl_analyzer.cpp:363-370
std::vector<size_t> found_gene_snarl(const std::vector<Qtl_data_t>& genes, size_t start, size_t end, size_t window);  // :471-491

// A column of values as the reference prints them, formatted on the device in one call (stoat_gpu_format_values):
// text(i) is what set_precision(values[i], codes[i]) returns.  The rows of a chunk are assembled from such columns
// into ONE buffer and written with one ostream::write — the reference formats every value through its own ostringstream
// and flushes the file after every row (src/writer.cpp:34,49,85), which is what caps its OpenMP scaling.
class TextColumn {
public:
    void format(stoat_gpu_ctx* ctx, const double* values, const uint8_t* codes, size_t n);
    void append_to(std::string& row, size_t i) const { row.append(slots.data() + 16 * i, lengths[i]); }

private:
    std::vector<char> slots;
    std::vector<uint8_t> lengths;
};

// Owner of the stoat_gpu_ctx handles; common state of the analyzers (what SnarlAnalyzer holds by reference,
// src/snarl_analyzer.hpp:75-92).  With one device the chunk is one batch call.  With several (`devices`), the chunk's snarl
// vector is cut into contiguous ranges balanced by path count, one host thread per device uploads the matrix, tests its
// range and formats its rows, and the fragments are written in snarl order: the snarl-sharded mode of SURVEY §8e behind
// the single-process C++ host (no collective; the reference's unit of parallelism is the snarl, :165-172).
class GpuSnarlAnalyzer {
public:
    GpuSnarlAnalyzer(const EdgeBySampleMatrix& edge_matrix, size_t min_individuals, size_t min_haplotypes,
                     double maf_threshold, const std::vector<int>& devices);
    virtual ~GpuSnarlAnalyzer();
    GpuSnarlAnalyzer(const GpuSnarlAnalyzer&) = delete;
    GpuSnarlAnalyzer& operator=(const GpuSnarlAnalyzer&) = delete;

    // Upload the chromosome's matrix (call after make_edge_matrix, src/snarl_analyzer.cpp:153) and test every snarl
    // of the chunk; rows of surviving snarls are written to outf in snarl order.  Returns the number of filtered
    // snarls (the sum of the reference's per-snarl bool, :169-172).
    size_t analyze_and_write_chunk(const std::vector<Snarl_data_t>& snarls, const std::string& chr, std::ostream& outf);
    virtual void write_header(std::ostream& outf) const = 0;
    stoat_gpu_timing last_timing() const;
    // wall-clock seconds of the last chunk's stages (slowest device): flatten, matrix upload + repack, phenotype,
    // batch call (H2D CSR, kernels, D2H), formatting + row assembly, the ostream write
    struct StageSeconds {
        double flatten = 0, set_matrix = 0, phenotype = 0, batch = 0, format = 0, write = 0;
    };
    StageSeconds last_stage_seconds() const { return stages; }
    size_t n_devices() const { return ctxs.size(); }

protected:
    virtual void upload_phenotype(stoat_gpu_ctx* ctx) = 0;
    // test snarls[0 .. n) on ctx and append their rows to `rows`; returns the number of filtered snarls
    virtual size_t run(stoat_gpu_ctx* ctx, const Snarl_data_t* snarls, size_t n, const FlatPaths& flat, const std::string& chr,
                       std::string& rows, StageSeconds& t) = 0;
    static void check(stoat_gpu_ctx* ctx, int rc, const char* what);  // throws std::runtime_error(last_error)

    const EdgeBySampleMatrix& edge_matrix;
    stoat_gpu_thresholds thr;
    std::vector<stoat_gpu_ctx*> ctxs;
    StageSeconds stages;
};

// BinarySnarlAnalyzer (src/snarl_analyzer.cpp:358-390): chi2 + Fisher per snarl
class GpuBinarySnarlAnalyzer : public GpuSnarlAnalyzer {
public:
    GpuBinarySnarlAnalyzer(const EdgeBySampleMatrix& edge_matrix, const std::vector<bool>& binary_phenotype,
                           size_t min_individuals, size_t min_haplotypes, double maf_threshold, int device = 0)
        : GpuBinarySnarlAnalyzer(edge_matrix, binary_phenotype, min_individuals, min_haplotypes, maf_threshold,
                                 std::vector<int>{device}) {}
    GpuBinarySnarlAnalyzer(const EdgeBySampleMatrix& edge_matrix, const std::vector<bool>& binary_phenotype,
                           size_t min_individuals, size_t min_haplotypes, double maf_threshold,
                           const std::vector<int>& devices);
    void write_header(std::ostream& outf) const override;  // src/writer.cpp:7-9

protected:
    void upload_phenotype(stoat_gpu_ctx* ctx) override;
    size_t run(stoat_gpu_ctx*, const Snarl_data_t*, size_t, const FlatPaths&, const std::string&, std::string&,
               StageSeconds&) override;
    const std::vector<bool>& binary_phenotype;  // false = control (g0), true = case (g1), src/arg_parser.cpp:20-95
};

// BinaryCovarSnarlAnalyzer (src/snarl_analyzer.cpp:392-428): logistic regression of the binary phenotype on path dosage.
// The reference never puts its covariates into the design (src/stats_test.cpp:59-62 shadows its own variables);
// use_covariates = false reproduces that, true appends them (stoat_gpu_binary_covar_batch).
class GpuBinaryCovarSnarlAnalyzer : public GpuSnarlAnalyzer {
public:
    GpuBinaryCovarSnarlAnalyzer(const EdgeBySampleMatrix& edge_matrix, const std::vector<bool>& binary_phenotype,
                                const std::vector<std::vector<double>>& covariate, size_t min_individuals,
                                size_t min_haplotypes, double maf_threshold, const std::vector<int>& devices = {0},
                                bool use_covariates = false);
    void write_header(std::ostream& outf) const override;  // src/writer.cpp:11-13

protected:
    void upload_phenotype(stoat_gpu_ctx* ctx) override;
    size_t run(stoat_gpu_ctx*, const Snarl_data_t*, size_t, const FlatPaths&, const std::string&, std::string&,
               StageSeconds&) override;
    const std::vector<bool>& binary_phenotype;
    const std::vector<std::vector<double>>& covariate;
    bool use_covariates;
};

// QuantitativeSnarlAnalyzer (src/snarl_analyzer.cpp:431-467): OLS of the phenotype on path dosage (+ covariates)
class GpuQuantitativeSnarlAnalyzer : public GpuSnarlAnalyzer {
public:
    GpuQuantitativeSnarlAnalyzer(const EdgeBySampleMatrix& edge_matrix, const std::vector<double>& quantitative_phenotype,
                                 const std::vector<std::vector<double>>& covariate, size_t min_individuals,
                                 size_t min_haplotypes, double maf_threshold, int device = 0)
        : GpuQuantitativeSnarlAnalyzer(edge_matrix, quantitative_phenotype, covariate, min_individuals, min_haplotypes,
                                       maf_threshold, std::vector<int>{device}) {}
    GpuQuantitativeSnarlAnalyzer(const EdgeBySampleMatrix& edge_matrix, const std::vector<double>& quantitative_phenotype,
                                 const std::vector<std::vector<double>>& covariate, size_t min_individuals,
                                 size_t min_haplotypes, double maf_threshold, const std::vector<int>& devices);
    void write_header(std::ostream& outf) const override;  // src/writer.cpp:15-17

protected:
    void upload_phenotype(stoat_gpu_ctx* ctx) override;
    size_t run(stoat_gpu_ctx*, const Snarl_data_t*, size_t, const FlatPaths&, const std::string&, std::string&,
               StageSeconds&) override;
    const std::vector<double>& quantitative_phenotype;
    const std::vector<std::vector<double>>& covariate;  // [sample][covariate]; empty = none
};

// EQTLSnarlAnalyzer (src/snarl_analyzer.cpp:493-539): one OLS per gene within windows_gene_threshold of the snarl
class GpuEQTLSnarlAnalyzer : public GpuSnarlAnalyzer {
public:
    GpuEQTLSnarlAnalyzer(const EdgeBySampleMatrix& edge_matrix,
                         const std::unordered_map<std::string, std::vector<Qtl_data_t>>& eqtl_map,
                         const std::vector<std::vector<double>>& covariate, size_t windows_gene_threshold,
                         size_t min_individuals, size_t min_haplotypes, double maf_threshold, int device = 0);
    void write_header(std::ostream& outf) const override;  // src/writer.cpp:19-21

protected:
    void upload_phenotype(stoat_gpu_ctx* ctx) override;
    size_t run(stoat_gpu_ctx*, const Snarl_data_t*, size_t, const FlatPaths&, const std::string&, std::string&,
               StageSeconds&) override;
    const std::unordered_map<std::string, std::vector<Qtl_data_t>>& eqtl_map;
    const std::vector<std::vector<double>>& covariate;
    size_t windows_gene_threshold;
};

}  // namespace stoat_gpu_shim
#endif  // STOAT_GPU_SHIM_HPP
