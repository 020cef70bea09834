// stoat_gpu_shim.cpp — see stoat_gpu_shim.hpp.  Host plumbing only: flattening, ABI calls, TSV rows.
#include "stoat_gpu_shim.hpp"

#include <algorithm>
#include <charconv>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <exception>
#include <thread>

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

namespace {
void flatten_range(const Snarl_data_t* snarls, size_t n_snarls, const EdgeBySampleMatrix& edge_matrix, FlatPaths& f) {
    f.snarl_path_off.reserve(n_snarls + 1);
    for (size_t si = 0; si < n_snarls; ++si) {
        const Snarl_data_t& s = snarls[si];
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
}
}  // namespace

// One hash lookup per edge (find_edge is const: the matrix is read-only during the chunk, as in the reference's OpenMP loop);
// large chunks are flattened by a few threads, each over a contiguous range of snarls, and the pieces are joined.
FlatPaths flatten_snarl_paths(const Snarl_data_t* snarls, size_t n_snarls, const EdgeBySampleMatrix& edge_matrix) {
    unsigned T = std::min(8u, std::max(1u, std::thread::hardware_concurrency() / 2));
    if (n_snarls < 65536) T = 1;
    FlatPaths f;
    if (T == 1) {
        flatten_range(snarls, n_snarls, edge_matrix, f);
    } else {
        std::vector<FlatPaths> part(T);
        std::vector<std::thread> th;
        for (unsigned t = 0; t < T; ++t) {
            const size_t a = n_snarls * t / T, b = n_snarls * (t + 1) / T;
            th.emplace_back([&, t, a, b] { flatten_range(snarls + a, b - a, edge_matrix, part[t]); });
        }
        for (std::thread& x : th) x.join();
        size_t np = 0, ne = 0;
        for (const FlatPaths& q : part) {
            np += q.path_edge_off.size() - 1;
            ne += q.edge_row.size();
        }
        f.snarl_path_off.reserve(n_snarls + 1);
        f.path_edge_off.reserve(np + 1);
        f.edge_row.reserve(ne + 1);
        for (const FlatPaths& q : part) {
            const uint64_t p0 = f.path_edge_off.size() - 1, e0 = f.edge_row.size();
            for (size_t i = 1; i < q.snarl_path_off.size(); ++i) f.snarl_path_off.push_back(q.snarl_path_off[i] + p0);
            for (size_t i = 1; i < q.path_edge_off.size(); ++i) f.path_edge_off.push_back(q.path_edge_off[i] + e0);
            f.edge_row.insert(f.edge_row.end(), q.edge_row.begin(), q.edge_row.end());
        }
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
double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
void put_uint(std::string& o, uint64_t v) {
    char buf[24];
    const auto r = std::to_chars(buf, buf + sizeof buf, v);
    o.append(buf, size_t(r.ptr - buf));
}
std::string snarl_name(const Snarl_data_t& s) {  // stoat::pairToString, src/snarl_data_t.cpp:181-185
    return std::to_string(s.snarl_ids.first) + "_" + std::to_string(s.snarl_ids.second);
}
void put_list(std::string& o, const uint32_t* v, size_t n) {  // stoat::vectorToString, src/utils.cpp:102-110
    for (size_t i = 0; i < n; ++i) {
        if (i) o += ',';
        put_uint(o, v[i]);
    }
}
// #CHR START_POS END_POS SNARL PATH_LENGTHS — the columns every writer starts with (src/writer.cpp:26-30)
void row_prefix(std::string& o, const std::string& chr, const Snarl_data_t& s) {
    o += chr;
    o += '\t';
    put_uint(o, s.start_positions);
    o += '\t';
    put_uint(o, s.end_positions);
    o += '\t';
    put_uint(o, s.snarl_ids.first);
    o += '_';
    put_uint(o, s.snarl_ids.second);
    o += '\t';
    for (size_t i = 0; i < s.type_variants.size(); ++i) {
        if (i) o += ',';
        o += s.type_variants[i];
    }
    o += '\t';
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

void TextColumn::format(stoat_gpu_ctx* ctx, const double* values, const uint8_t* codes, size_t n) {
    slots.assign(16 * std::max<size_t>(n, 1), 0);
    lengths.assign(std::max<size_t>(n, 1), 0);
    const int rc = stoat_gpu_format_values(ctx, values, codes, n, slots.data(), lengths.data());
    if (rc != STOAT_GPU_OK) throw std::runtime_error(std::string("format_values: ") + stoat_gpu_last_error(ctx));
}

// ---- base ------------------------------------------------------------------------------------------------------------
GpuSnarlAnalyzer::GpuSnarlAnalyzer(const EdgeBySampleMatrix& m, size_t min_individuals, size_t min_haplotypes,
                                   double maf_threshold, const std::vector<int>& devices)
    : edge_matrix(m) {
    thr.min_individuals = min_individuals;
    thr.min_haplotypes = min_haplotypes;
    thr.maf_threshold = maf_threshold;
    if (devices.empty()) throw std::runtime_error("GpuSnarlAnalyzer: no device");
    for (int d : devices) {
        stoat_gpu_ctx* c = nullptr;
        if (stoat_gpu_create(&c, d) != STOAT_GPU_OK) {
            const std::string msg = stoat_gpu_last_error(nullptr);
            for (stoat_gpu_ctx* x : ctxs) stoat_gpu_destroy(x);
            throw std::runtime_error("stoat_gpu_create: " + msg);
        }
        ctxs.push_back(c);
    }
}

GpuSnarlAnalyzer::~GpuSnarlAnalyzer() {
    for (stoat_gpu_ctx* c : ctxs) stoat_gpu_destroy(c);
}

void GpuSnarlAnalyzer::check(stoat_gpu_ctx* ctx, int rc, const char* what) {
    if (rc != STOAT_GPU_OK) throw std::runtime_error(std::string(what) + ": " + stoat_gpu_last_error(ctx));
}

stoat_gpu_timing GpuSnarlAnalyzer::last_timing() const {
    stoat_gpu_timing t{};
    stoat_gpu_last_timing(ctxs[0], &t);
    return t;
}

size_t GpuSnarlAnalyzer::analyze_and_write_chunk(const std::vector<Snarl_data_t>& snarls, const std::string& chr,
                                                 std::ostream& outf) {
    stages = StageSeconds{};
    if (snarls.empty()) return 0;
    const size_t D = std::min(ctxs.size(), snarls.size());
    // contiguous ranges with about the same number of paths each (the work of a snarl grows with its paths)
    std::vector<size_t> cut(D + 1, snarls.size());
    cut[0] = 0;
    if (D > 1) {
        size_t total = 0;
        for (const Snarl_data_t& s : snarls) total += s.snarl_paths.size() + 1;
        size_t acc = 0, next = 1;
        for (size_t i = 0; i < snarls.size() && next < D; ++i) {
            acc += snarls[i].snarl_paths.size() + 1;
            if (acc * D >= total * next) cut[next++] = i + 1;
        }
    }
    std::vector<std::string> rows(D);
    std::vector<size_t> n_filtered(D, 0);
    std::vector<StageSeconds> t(D);
    std::vector<std::exception_ptr> err(D);
    auto work = [&](size_t d) {
        try {
            stoat_gpu_ctx* ctx = ctxs[d];
            const size_t s0 = cut[d], s1 = cut[d + 1];
            if (s1 <= s0) return;
            double t0 = now_s();
            check(ctx, stoat_gpu_set_matrix(ctx, edge_matrix.data().data(), edge_matrix.rows(), edge_matrix.cols()), "set_matrix");
            t[d].set_matrix = now_s() - t0;
            t0 = now_s();
            upload_phenotype(ctx);
            t[d].phenotype = now_s() - t0;
            t0 = now_s();
            const FlatPaths flat = flatten_snarl_paths(snarls.data() + s0, s1 - s0, edge_matrix);
            t[d].flatten = now_s() - t0;
            n_filtered[d] = run(ctx, snarls.data() + s0, s1 - s0, flat, chr, rows[d], t[d]);
        } catch (...) {
            err[d] = std::current_exception();
        }
    };
    if (D == 1) {
        work(0);
    } else {
        std::vector<std::thread> th;
        for (size_t d = 0; d < D; ++d) th.emplace_back(work, d);
        for (std::thread& x : th) x.join();
    }
    for (const std::exception_ptr& e : err)
        if (e) std::rethrow_exception(e);
    const double t0 = now_s();
    size_t total_filtered = 0;
    for (size_t d = 0; d < D; ++d) {
        outf.write(rows[d].data(), std::streamsize(rows[d].size()));  // fragments in snarl order, one write each
        total_filtered += n_filtered[d];
        stages.flatten = std::max(stages.flatten, t[d].flatten);
        stages.set_matrix = std::max(stages.set_matrix, t[d].set_matrix);
        stages.phenotype = std::max(stages.phenotype, t[d].phenotype);
        stages.batch = std::max(stages.batch, t[d].batch);
        stages.format = std::max(stages.format, t[d].format);
    }
    stages.write = now_s() - t0;
    return total_filtered;
}

// ---- binary ----------------------------------------------------------------------------------------------------------
GpuBinarySnarlAnalyzer::GpuBinarySnarlAnalyzer(const EdgeBySampleMatrix& m, const std::vector<bool>& pheno,
                                               size_t min_individuals, size_t min_haplotypes, double maf_threshold,
                                               const std::vector<int>& devices)
    : GpuSnarlAnalyzer(m, min_individuals, min_haplotypes, maf_threshold, devices), binary_phenotype(pheno) {}

void GpuBinarySnarlAnalyzer::write_header(std::ostream& outf) const {
    outf << "#CHR\tSTART_POS\tEND_POS\tSNARL\tPATH_LENGTHS\tP_FISHER\tP_CHI2\tGROUP_PATHS\tDEPTH\n";
}

namespace {
void upload_binary(stoat_gpu_ctx* ctx, const std::vector<bool>& pheno) {
    std::vector<uint8_t> is_case(pheno.size());
    for (size_t i = 0; i < is_case.size(); ++i) is_case[i] = pheno[i] ? 1 : 0;
    const int rc = stoat_gpu_set_binary_phenotype(ctx, is_case.data(), is_case.size());
    if (rc != STOAT_GPU_OK) throw std::runtime_error(std::string("set_binary_phenotype: ") + stoat_gpu_last_error(ctx));
}
}  // namespace

void GpuBinarySnarlAnalyzer::upload_phenotype(stoat_gpu_ctx* ctx) { upload_binary(ctx, binary_phenotype); }

size_t GpuBinarySnarlAnalyzer::run(stoat_gpu_ctx* ctx, const Snarl_data_t* snarls, size_t S, const FlatPaths& flat,
                                   const std::string& chr, std::string& rows, StageSeconds& t) {
    const size_t P = flat.path_edge_off.size() - 1;
    std::vector<uint32_t> g0(P), g1(P);
    std::vector<uint64_t> total(S);
    std::vector<uint8_t> filtered(S), chi2_code(S), fisher_code(S);
    std::vector<double> p_chi2(S), p_fisher(S);
    stoat_gpu_binary_out out{g0.data(), g1.data(), total.data(), filtered.data(), p_chi2.data(), p_fisher.data(),
                             chi2_code.data(), fisher_code.data()};
    const stoat_gpu_paths pv = flat.view();
    double t0 = now_s();
    check(ctx, stoat_gpu_binary_batch(ctx, &pv, &thr, &out), "binary_batch");
    t.batch = now_s() - t0;
    t0 = now_s();
    TextColumn fisher, chi2;
    fisher.format(ctx, p_fisher.data(), fisher_code.data(), S);
    chi2.format(ctx, p_chi2.data(), chi2_code.data(), S);
    size_t n_filtered = 0;
    rows.reserve(rows.size() + S * 72);
    for (size_t s = 0; s < S; ++s) {
        if (filtered[s]) {
            ++n_filtered;
            continue;
        }
        row_prefix(rows, chr, snarls[s]);  // stoat::write_binary, src/writer.cpp:23-35
        fisher.append_to(rows, s);
        rows += '\t';
        chi2.append_to(rows, s);
        rows += '\t';
        // remove_empty_columns_binary_table (src/snarl_analyzer.cpp:686-702): GROUP_PATHS prints the kept columns
        bool first = true;
        for (uint64_t p = flat.snarl_path_off[s]; p < flat.snarl_path_off[s + 1]; ++p)
            if (g0[p] + g1[p] != 0) {
                if (!first) rows += ',';
                first = false;
                put_uint(rows, g0[p]);  // format_group_paths, src/binary_table.cpp:6-17
                rows += ':';
                put_uint(rows, g1[p]);
            }
        rows += '\t';
        put_uint(rows, snarls[s].depth);
        rows += '\n';
    }
    t.format = now_s() - t0;
    return n_filtered;
}

// ---- binary with covariates (logistic regression) --------------------------------------------------------------------
GpuBinaryCovarSnarlAnalyzer::GpuBinaryCovarSnarlAnalyzer(const EdgeBySampleMatrix& m, const std::vector<bool>& pheno,
                                                         const std::vector<std::vector<double>>& covar,
                                                         size_t min_individuals, size_t min_haplotypes, double maf_threshold,
                                                         const std::vector<int>& devices, bool use_covar)
    : GpuSnarlAnalyzer(m, min_individuals, min_haplotypes, maf_threshold, devices), binary_phenotype(pheno),
      covariate(covar), use_covariates(use_covar) {}

void GpuBinaryCovarSnarlAnalyzer::write_header(std::ostream& outf) const {
    outf << "#CHR\tSTART_POS\tEND_POS\tSNARL\tPATH_LENGTHS\tP\tBETA\tSE\tALLELE_PATHS\tDEPTH\n";
}

void GpuBinaryCovarSnarlAnalyzer::upload_phenotype(stoat_gpu_ctx* ctx) {
    upload_binary(ctx, binary_phenotype);
    size_t n_covar = 0;
    const std::vector<double> flat = flatten_covariates(covariate, n_covar);
    check(ctx, stoat_gpu_set_covariates(ctx, n_covar ? flat.data() : nullptr, covariate.size(), n_covar), "set_covariates");
}

namespace {
void check_ols_code(uint8_t code, const Snarl_data_t& s) {
    if (code == STOAT_P_OVERLIMIT)
        throw std::runtime_error("snarl " + snarl_name(s) + " exceeds STOAT_GPU_MAX_PATHS_PER_SNARL / "
                                 "STOAT_GPU_MAX_OLS_FEATURES: run the per-snarl host analyzer for it");
}
}  // namespace

size_t GpuBinaryCovarSnarlAnalyzer::run(stoat_gpu_ctx* ctx, const Snarl_data_t* snarls, size_t S, const FlatPaths& flat,
                                        const std::string& chr, std::string& rows, StageSeconds& t) {
    const size_t P = flat.path_edge_off.size() - 1;
    std::vector<uint32_t> allele(P), n_used(S);
    std::vector<uint8_t> filtered(S), code(S);
    std::vector<double> p(S), beta(S), se(S);
    stoat_gpu_binary_covar_out out{allele.data(), filtered.data(), n_used.data(), p.data(), beta.data(), se.data(), code.data()};
    const stoat_gpu_paths pv = flat.view();
    double t0 = now_s();
    check(ctx, stoat_gpu_binary_covar_batch(ctx, &pv, &thr, use_covariates ? 1 : 0, &out), "binary_covar_batch");
    t.batch = now_s() - t0;
    t0 = now_s();
    // "NA" for all three when the fit failed (src/stats_test.cpp:107,120): the code column says so
    TextColumn tp, tb, ts;
    tp.format(ctx, p.data(), code.data(), S);
    tb.format(ctx, beta.data(), code.data(), S);
    ts.format(ctx, se.data(), code.data(), S);
    size_t n_filtered = 0;
    for (size_t s = 0; s < S; ++s) {
        if (filtered[s]) {
            ++n_filtered;
            continue;
        }
        check_ols_code(code[s], snarls[s]);
        const uint64_t p0 = flat.snarl_path_off[s], np = flat.snarl_path_off[s + 1] - p0;
        row_prefix(rows, chr, snarls[s]);  // stoat::write_binary_covar, src/writer.cpp:37-50
        tp.append_to(rows, s);
        rows += '\t';
        tb.append_to(rows, s);
        rows += '\t';
        ts.append_to(rows, s);
        rows += '\t';
        put_list(rows, allele.data() + p0, np);
        rows += '\t';
        put_uint(rows, snarls[s].depth);
        rows += '\n';
    }
    t.format = now_s() - t0;
    return n_filtered;
}

// ---- quantitative ----------------------------------------------------------------------------------------------------
GpuQuantitativeSnarlAnalyzer::GpuQuantitativeSnarlAnalyzer(const EdgeBySampleMatrix& m, const std::vector<double>& pheno,
                                                           const std::vector<std::vector<double>>& covar,
                                                           size_t min_individuals, size_t min_haplotypes,
                                                           double maf_threshold, const std::vector<int>& devices)
    : GpuSnarlAnalyzer(m, min_individuals, min_haplotypes, maf_threshold, devices), quantitative_phenotype(pheno),
      covariate(covar) {}

void GpuQuantitativeSnarlAnalyzer::write_header(std::ostream& outf) const {
    outf << "#CHR\tSTART_POS\tEND_POS\tSNARL\tPATH_LENGTHS\tP\tRSQUARE\tBETA\tSE\tALLELE_PATHS\tDEPTH\n";
}

void GpuQuantitativeSnarlAnalyzer::upload_phenotype(stoat_gpu_ctx* ctx) {
    check(ctx, stoat_gpu_set_quantitative_phenotype(ctx, quantitative_phenotype.data(), quantitative_phenotype.size()),
          "set_quantitative_phenotype");
    size_t n_covar = 0;
    const std::vector<double> flat = flatten_covariates(covariate, n_covar);
    check(ctx, stoat_gpu_set_covariates(ctx, n_covar ? flat.data() : nullptr, covariate.size(), n_covar), "set_covariates");
}

size_t GpuQuantitativeSnarlAnalyzer::run(stoat_gpu_ctx* ctx, const Snarl_data_t* snarls, size_t S, const FlatPaths& flat,
                                         const std::string& chr, std::string& rows, StageSeconds& t) {
    const size_t P = flat.path_edge_off.size() - 1;
    std::vector<uint32_t> allele(P), n_used(S);
    std::vector<uint8_t> filtered(S), code(S);
    std::vector<double> p(S), r2(S), beta(S), se(S);
    stoat_gpu_quant_out out{allele.data(), filtered.data(), n_used.data(), p.data(), r2.data(), beta.data(), se.data(),
                            code.data()};
    const stoat_gpu_paths pv = flat.view();
    double t0 = now_s();
    check(ctx, stoat_gpu_quantitative_batch(ctx, &pv, &thr, &out), "quantitative_batch");
    t.batch = now_s() - t0;
    t0 = now_s();
    TextColumn tp, tr, tb, ts;
    tp.format(ctx, p.data(), nullptr, S);
    tr.format(ctx, r2.data(), nullptr, S);
    tb.format(ctx, beta.data(), nullptr, S);
    ts.format(ctx, se.data(), nullptr, S);
    size_t n_filtered = 0;
    rows.reserve(rows.size() + S * 96);
    for (size_t s = 0; s < S; ++s) {
        if (filtered[s]) {
            ++n_filtered;
            continue;
        }
        check_ols_code(code[s], snarls[s]);
        const uint64_t p0 = flat.snarl_path_off[s], np = flat.snarl_path_off[s + 1] - p0;
        row_prefix(rows, chr, snarls[s]);  // stoat::write_quantitative, src/writer.cpp:53-68
        tp.append_to(rows, s);
        rows += '\t';
        tr.append_to(rows, s);
        rows += '\t';
        tb.append_to(rows, s);
        rows += '\t';
        ts.append_to(rows, s);
        rows += '\t';
        put_list(rows, allele.data() + p0, np);
        rows += '\t';
        put_uint(rows, snarls[s].depth);
        rows += '\n';
    }
    t.format = now_s() - t0;
    return n_filtered;
}

// ---- eQTL ------------------------------------------------------------------------------------------------------------
GpuEQTLSnarlAnalyzer::GpuEQTLSnarlAnalyzer(const EdgeBySampleMatrix& m,
                                           const std::unordered_map<std::string, std::vector<Qtl_data_t>>& map,
                                           const std::vector<std::vector<double>>& covar, size_t window,
                                           size_t min_individuals, size_t min_haplotypes, double maf_threshold, int device)
    : GpuSnarlAnalyzer(m, min_individuals, min_haplotypes, maf_threshold, std::vector<int>{device}), eqtl_map(map),
      covariate(covar), windows_gene_threshold(window) {}

void GpuEQTLSnarlAnalyzer::write_header(std::ostream& outf) const {
    outf << "#CHR\tSTART_POS\tEND_POS\tSNARL\tPATH_LENGTHS\tGENE\tP\tRSQUARE\tBETA\tSE\tALLELE_PATHS\tDEPTH\n";
}

void GpuEQTLSnarlAnalyzer::upload_phenotype(stoat_gpu_ctx* ctx) {
    size_t n_covar = 0;
    const std::vector<double> flat = flatten_covariates(covariate, n_covar);
    check(ctx, stoat_gpu_set_covariates(ctx, n_covar ? flat.data() : nullptr, covariate.size(), n_covar), "set_covariates");
}

size_t GpuEQTLSnarlAnalyzer::run(stoat_gpu_ctx* ctx, const Snarl_data_t* snarls, size_t S, const FlatPaths& flat,
                                 const std::string& chr, std::string& rows, StageSeconds& t) {
    const std::vector<Qtl_data_t>& genes = eqtl_map.at(chr);  // throws like the reference's .at() (:496)
    const size_t P = flat.path_edge_off.size() - 1, G = genes.size();
    const size_t n = edge_matrix.cols() / 2;
    // gene windows on the host (found_gene_snarl, :471-491) -> CSR
    std::vector<uint64_t> off{0};
    std::vector<uint32_t> idx;
    for (size_t s = 0; s < S; ++s) {
        for (size_t g : found_gene_snarl(genes, snarls[s].start_positions, snarls[s].end_positions, windows_gene_threshold))
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
    double t0 = now_s();
    check(ctx, stoat_gpu_eqtl_batch(ctx, &pv, &gv, &thr, &out), "eqtl_batch");
    t.batch = now_s() - t0;
    t0 = now_s();
    TextColumn tp, tr, tb, ts;
    tp.format(ctx, p.data(), nullptr, n_pairs);
    tr.format(ctx, r2.data(), nullptr, n_pairs);
    tb.format(ctx, beta.data(), nullptr, n_pairs);
    ts.format(ctx, se.data(), nullptr, n_pairs);
    size_t n_filtered = 0;
    for (size_t s = 0; s < S; ++s) {
        if (filtered[s]) {
            ++n_filtered;
            continue;
        }
        const uint64_t p0 = flat.snarl_path_off[s], np = flat.snarl_path_off[s + 1] - p0;
        for (uint64_t k = off[s]; k < off[s + 1]; ++k) {
            check_ols_code(code[k], snarls[s]);
            row_prefix(rows, chr, snarls[s]);  // stoat::write_eqtl, src/writer.cpp:70-87
            rows += genes[idx[k]].geneName;
            rows += '\t';
            tp.append_to(rows, k);
            rows += '\t';
            tr.append_to(rows, k);
            rows += '\t';
            tb.append_to(rows, k);
            rows += '\t';
            ts.append_to(rows, k);
            rows += '\t';
            put_list(rows, allele.data() + p0, np);
            rows += '\t';
            put_uint(rows, snarls[s].depth);
            rows += '\n';
        }
    }
    t.format = now_s() - t0;
    return n_filtered;
}

}  // namespace stoat_gpu_shim
