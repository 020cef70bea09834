// shim_selftest.cpp — drives the C++ host shim the way the reference's main_stoat drives its analyzers
// (src/subcommand/vcf.cpp:363-395): build the EdgeBySampleMatrix with push_matrix, construct the analyzer, one
// analyze_and_write_chunk per chromosome, rows into a TSV.  Used by tests/test_gpu_shim.py (needs a B200).
//
//   shim_selftest golden <binary_table_vcf.golden.tsv> <binary_phenotype.tsv> <out.tsv>
//       every golden row of the reference's expected output becomes a snarl whose paths are walked by exactly the
//       haplotypes its GROUP_PATHS column counts; the written TSV must equal the golden file row for row.
//   shim_selftest bench <n_snarls> <n_samples> <devices, e.g. 0 or 0,1> <out.json>
//       a C3-shaped chunk (biallelic snarls, random matrix bytes) through GpuBinarySnarlAnalyzer: seconds per stage
//       (flatten / matrix upload + repack / batch / format / write) and the same rows written the reference's way
//       (an ostringstream per value, std::endl per row: src/utils.cpp:5-15, src/writer.cpp:23-35) for comparison.
//   shim_selftest synth <seed> <n_snarls> <n_samples> <out_dir>
//       random haplotype walks; writes binary.tsv / quantitative.tsv / quantitative_covar.tsv / eqtl.tsv and dumps the
//       problem (matrix bytes, CSR, phenotypes, expression, gene windows) as raw arrays for the Python side to feed
//       the oracle.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>

#include "stoat_gpu_shim.hpp"

using namespace stoat_gpu_shim;

namespace {

std::vector<std::string> split(const std::string& s, char sep) {
    std::vector<std::string> out;
    std::string item;
    std::istringstream is(s);
    while (std::getline(is, item, sep)) out.push_back(item);
    return out;
}

template <typename T>
void dump(const std::string& path, const std::vector<T>& v) {
    std::ofstream f(path, std::ios::binary);
    f.write(reinterpret_cast<const char*>(v.data()), std::streamsize(v.size() * sizeof(T)));
}

// a snarl record to be filled in: the reference's Snarl_data_t has no default constructor (src/snarl_data_t.hpp:98-118)
Snarl_data_t blank_snarl() {
#ifdef STOAT_GPU_SHIM_USE_REFERENCE_TYPES
    return Snarl_data_t(handlegraph::net_handle_t(), std::pair<size_t, size_t>(0, 0), std::vector<Path_traversal_t>(), 0, 0,
                        std::vector<std::string>(), 0);
#else
    return Snarl_data_t();
#endif
}

Path_traversal_t make_path(std::initializer_list<size_t> nodes) {
    Path_traversal_t p;
    for (size_t n : nodes) p.add_node_traversal_t(Node_traversal_t(n, false));
    return p;
}

// a haplotype walking `path` sets the bit of every edge of the path (make_edge_matrix, src/snarl_analyzer.cpp:190-260)
void walk(EdgeBySampleMatrix& m, const Path_traversal_t& path, size_t hap) {
    const auto& nodes = path.get_paths();
    for (size_t i = 0; i + 1 < nodes.size(); ++i) {
        if (nodes[i].get_node_id() == 0 || nodes[i + 1].get_node_id() == 0) continue;
        m.push_matrix(Edge_t(nodes[i], nodes[i + 1]), hap);
    }
}

int run_golden(const char* golden_tsv, const char* pheno_tsv, const char* out_tsv) {
    // phenotype: last column PHENO with 1 = control, 2 = case (parse_binary_pheno, src/arg_parser.cpp:20-95)
    std::vector<bool> pheno;
    {
        std::ifstream f(pheno_tsv);
        std::string line;
        std::getline(f, line);
        while (std::getline(f, line)) {
            const auto c = split(line, '\t');
            if (c.size() >= 2) pheno.push_back(c.back() == "2");
        }
    }
    const size_t n = pheno.size(), H = 2 * n;
    std::vector<size_t> ctrl, cases;
    for (size_t h = 0; h < H; ++h) (pheno[h / 2] ? cases : ctrl).push_back(h);
    std::vector<std::string> names(n);
    EdgeBySampleMatrix matrix(names, 4, H);
    std::vector<Snarl_data_t> snarls;
    std::string chr;
    std::ifstream f(golden_tsv);
    std::string line;
    std::getline(f, line);
    size_t next_node = 1000000;
    while (std::getline(f, line)) {
        const auto c = split(line, '\t');
        if (c.size() < 9) continue;
        chr = c[0];
        Snarl_data_t s = blank_snarl();
        s.start_positions = std::stoull(c[1]);
        s.end_positions = std::stoull(c[2]);
        const auto ids = split(c[3], '_');
        s.snarl_ids = {std::stoull(ids[0]), std::stoull(ids[1])};
        s.type_variants = split(c[4], ',');
        s.depth = std::stoull(c[8]);
        size_t c0 = 0, c1 = 0;
        for (const std::string& col : split(c[7], ',')) {  // GROUP_PATHS "g0:g1,g0:g1"
            const auto g = split(col, ':');
            const size_t g0 = std::stoull(g[0]), g1 = std::stoull(g[1]);
            const Path_traversal_t path = make_path({s.snarl_ids.second, next_node++, s.snarl_ids.first});
            for (size_t k = 0; k < g0; ++k) walk(matrix, path, ctrl.at(c0++));
            for (size_t k = 0; k < g1; ++k) walk(matrix, path, cases.at(c1++));
            s.snarl_paths.push_back(path);
        }
        snarls.push_back(std::move(s));
    }
    GpuBinarySnarlAnalyzer analyzer(matrix, pheno, 3, 5, 0.05);  // defaults of `stoat vcf` (src/subcommand/vcf.cpp:61-73)
    std::ofstream out(out_tsv);
    analyzer.write_header(out);
    const size_t filtered = analyzer.analyze_and_write_chunk(snarls, chr, out);
    std::printf("golden: %zu snarls, %zu filtered, %zu matrix rows\n", snarls.size(), filtered, matrix.rows());
    return 0;
}

struct Rng {  // splitmix64
    uint64_t s;
    uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    double uniform() { return double(next() >> 11) * (1.0 / 9007199254740992.0); }
    double normal() {
        double a = 0;
        for (int i = 0; i < 12; ++i) a += uniform();
        return a - 6.0;
    }
};

int run_synth(uint64_t seed, size_t n_snarls, size_t n_samples, const std::string& dir) {
    Rng rng{seed};
    const size_t H = 2 * n_samples;
    std::vector<std::string> names(n_samples);
    EdgeBySampleMatrix matrix(names, 4, H);
    std::vector<Snarl_data_t> snarls;
    size_t node = 10;
    for (size_t s = 0; s < n_snarls; ++s) {
        Snarl_data_t sd = blank_snarl();
        const size_t a = node++, d = node++;
        sd.snarl_ids = {d, a};
        sd.start_positions = 1000 * s + 17;
        sd.end_positions = sd.start_positions + rng.next() % 300;
        sd.depth = 1 + rng.next() % 3;
        const size_t np = 2 + rng.next() % 3;
        for (size_t j = 0; j < np; ++j) {
            sd.type_variants.push_back(std::to_string(rng.next() % 50));
            const uint64_t kind = rng.next() % 20;
            if (kind == 0) sd.snarl_paths.push_back(make_path({a, 0, d}));              // complex: no checked edge
            else if (kind == 1) sd.snarl_paths.push_back(make_path({a, node++, node++, d}));  // three edges
            else sd.snarl_paths.push_back(make_path({a, node++, d}));
        }
        if (rng.next() % 15 == 0) sd.snarl_paths.push_back(make_path({a, node++, d}));  // a path nobody walks
        const double f = rng.uniform();
        for (size_t h = 0; h < H; ++h) {
            if (rng.uniform() < 0.02) continue;  // haplotype on no path
            const size_t j = rng.uniform() < f ? 0 : 1 + rng.next() % (np - 1);
            walk(matrix, sd.snarl_paths[j], h);
        }
        snarls.push_back(std::move(sd));
    }
    std::vector<bool> is_case(n_samples);
    std::vector<double> y(n_samples);
    std::vector<std::vector<double>> covar(n_samples, std::vector<double>(3));
    for (size_t i = 0; i < n_samples; ++i) {
        is_case[i] = rng.uniform() < 0.5;
        y[i] = 3.0 * rng.normal() + 10.0;
        for (double& c : covar[i]) c = rng.normal();
    }
    const size_t n_genes = 12;
    std::unordered_map<std::string, std::vector<Qtl_data_t>> eqtl_map;
    std::vector<Qtl_data_t>& genes = eqtl_map["chr7"];
    for (size_t g = 0; g < n_genes; ++g) {
        Qtl_data_t q;
        q.geneName = "GENE" + std::to_string(g);
        q.start_pos = 1000 * (g * n_snarls / n_genes);
        q.end_pos = q.start_pos + 2500;
        const double base = 8.0 * rng.uniform();
        for (size_t i = 0; i < n_samples; ++i) q.sampleExpresion.push_back(base * (1.0 + (2.0 * rng.uniform() - 1.0)));
        genes.push_back(std::move(q));
    }
    const std::vector<std::vector<double>> no_covar;
    const size_t window = 3000;
    {
        GpuBinarySnarlAnalyzer a(matrix, is_case, 3, 5, 0.05);
        std::ofstream out(dir + "/binary.tsv");
        a.write_header(out);
        a.analyze_and_write_chunk(snarls, "chr7", out);
    }
    {
        GpuQuantitativeSnarlAnalyzer a(matrix, y, no_covar, 3, 5, 0.05);
        std::ofstream out(dir + "/quantitative.tsv");
        a.write_header(out);
        a.analyze_and_write_chunk(snarls, "chr7", out);
    }
    {
        GpuQuantitativeSnarlAnalyzer a(matrix, y, covar, 3, 5, 0.05);
        std::ofstream out(dir + "/quantitative_covar.tsv");
        a.write_header(out);
        a.analyze_and_write_chunk(snarls, "chr7", out);
    }
    {
        // logistic regression (BinaryCovarSnarlAnalyzer); covariates passed and, as in the reference, without effect
        GpuBinaryCovarSnarlAnalyzer a(matrix, is_case, covar, 3, 5, 0.05);
        std::ofstream out(dir + "/binary_covar.tsv");
        a.write_header(out);
        a.analyze_and_write_chunk(snarls, "chr7", out);
    }
    {
        // the same chunk cut over several contexts on several host threads (STOAT_SHIM_DEVICES, default "0,0,0": three
        // contexts on GPU 0, so the path also runs on a one-GPU box); the rows must equal binary.tsv
        std::vector<int> devs;
        const char* e = std::getenv("STOAT_SHIM_DEVICES");
        for (const std::string& d : split(e ? e : "0,0,0", ',')) devs.push_back(std::atoi(d.c_str()));
        GpuBinarySnarlAnalyzer a(matrix, is_case, 3, 5, 0.05, devs);
        std::ofstream out(dir + "/binary_multi.tsv");
        a.write_header(out);
        a.analyze_and_write_chunk(snarls, "chr7", out);
        GpuQuantitativeSnarlAnalyzer q(matrix, y, no_covar, 3, 5, 0.05, devs);
        std::ofstream outq(dir + "/quantitative_multi.tsv");
        q.write_header(outq);
        q.analyze_and_write_chunk(snarls, "chr7", outq);
    }
    {
        GpuEQTLSnarlAnalyzer a(matrix, eqtl_map, no_covar, window, 3, 5, 0.05);
        std::ofstream out(dir + "/eqtl.tsv");
        a.write_header(out);
        a.analyze_and_write_chunk(snarls, "chr7", out);
    }
    // the problem, for the oracle
    const FlatPaths flat = flatten_snarl_paths(snarls, matrix);
    std::vector<uint8_t> bits(matrix.data().begin(), matrix.data().begin() + (matrix.rows() * H + 7) / 8);
    dump(dir + "/bits.u8", bits);
    dump(dir + "/snarl_path_off.u64", flat.snarl_path_off);
    dump(dir + "/path_edge_off.u64", flat.path_edge_off);
    dump(dir + "/edge_row.u32", flat.edge_row);
    std::vector<uint8_t> case_u8(is_case.begin(), is_case.end());
    dump(dir + "/is_case.u8", case_u8);
    dump(dir + "/y.f64", y);
    std::vector<double> cflat, expr;
    for (const auto& r : covar) cflat.insert(cflat.end(), r.begin(), r.end());
    dump(dir + "/covar.f64", cflat);
    std::vector<uint64_t> goff{0};
    std::vector<uint32_t> gidx;
    for (const Snarl_data_t& s : snarls) {
        for (size_t g : found_gene_snarl(genes, s.start_positions, s.end_positions, window)) gidx.push_back(uint32_t(g));
        goff.push_back(gidx.size());
    }
    for (const auto& g : genes) expr.insert(expr.end(), g.sampleExpresion.begin(), g.sampleExpresion.end());
    dump(dir + "/expression.f64", expr);
    dump(dir + "/gene_off.u64", goff);
    dump(dir + "/gene_idx.u32", gidx);
    std::ofstream meta(dir + "/meta.txt");
    meta << matrix.rows() << " " << H << " " << n_snarls << " " << n_genes << "\n";
    for (const Snarl_data_t& s : snarls)
        meta << s.start_positions << "\t" << s.end_positions << "\t" << s.snarl_ids.first << "_" << s.snarl_ids.second << "\t"
             << type_variants_string(s) << "\t" << s.depth << "\n";
    std::printf("synth: %zu snarls, %zu matrix rows, %zu gene pairs\n", n_snarls, matrix.rows(), gidx.size());
    return 0;
}


// ---- bench ---------------------------------------------------------------------------------------------------------------
// EdgeBySampleMatrix keeps its storage protected; a derived class may fill it (what make_edge_matrix does bit by bit)
struct FilledMatrix : EdgeBySampleMatrix {
    FilledMatrix(const std::vector<std::string>& names, size_t rows, size_t cols) : EdgeBySampleMatrix(names, rows, cols) {}
    void fill(Rng& rng, const std::vector<Edge_t>& edges) {
        uint64_t* w = reinterpret_cast<uint64_t*>(matrix_1D.data());
        for (size_t i = 0; i < matrix_1D.size() / 8; ++i) w[i] = rng.next();
        row_header.reserve(edges.size());
        for (size_t r = 0; r < edges.size(); ++r) row_header.emplace(edges[r], r);
    }
};

double wall() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

std::string reference_set_precision(double value) {  // the reference's own way, src/utils.cpp:5-15
    std::ostringstream oss;
    if (std::abs(value) < 1e-1 && value != 0.0) oss << std::scientific << std::setprecision(4);
    else oss << std::defaultfloat << std::setprecision(4);
    oss << value;
    return oss.str();
}

int run_bench(size_t n_snarls, size_t n_samples, const std::string& devices, const char* out_json) {
    Rng rng{7};
    const size_t H = 2 * n_samples;
    std::vector<std::string> names(n_samples);
    std::vector<Snarl_data_t> snarls(n_snarls, blank_snarl());
    std::vector<Edge_t> edges;
    edges.reserve(4 * n_snarls);
    size_t node = 10;
    for (size_t s = 0; s < n_snarls; ++s) {
        Snarl_data_t& sd = snarls[s];
        const size_t a = node++, b = node++, c = node++, d = node++;
        sd.snarl_ids = {d, a};
        sd.start_positions = 100 * s + 17;
        sd.end_positions = sd.start_positions + 1;
        sd.depth = 1;
        sd.type_variants = {"1", "1"};
        sd.snarl_paths.push_back(make_path({a, b, d}));
        sd.snarl_paths.push_back(make_path({a, c, d}));
        edges.emplace_back(Node_traversal_t(a, false), Node_traversal_t(b, false));
        edges.emplace_back(Node_traversal_t(b, false), Node_traversal_t(d, false));
        edges.emplace_back(Node_traversal_t(a, false), Node_traversal_t(c, false));
        edges.emplace_back(Node_traversal_t(c, false), Node_traversal_t(d, false));
    }
    FilledMatrix matrix(names, edges.size(), H);
    matrix.fill(rng, edges);
    std::vector<bool> is_case(n_samples);
    for (size_t i = 0; i < n_samples; ++i) is_case[i] = i >= n_samples / 2;
    std::vector<int> devs;
    for (const std::string& d : split(devices, ',')) devs.push_back(std::atoi(d.c_str()));
    GpuBinarySnarlAnalyzer an(matrix, is_case, 3, 5, 0.05, devs);
    const std::string tsv = "/tmp/shim_bench.tsv";
    GpuSnarlAnalyzer::StageSeconds st;
    double chunk_s = 0;
    size_t filtered = 0;
    for (int rep = 0; rep < 2; ++rep) {  // the second pass is the measured one (allocations, page faults)
        std::ofstream out(tsv);
        an.write_header(out);
        const double t0 = wall();
        filtered = an.analyze_and_write_chunk(snarls, "chr1", out);
        out.flush();
        chunk_s = wall() - t0;
        st = an.last_stage_seconds();
    }
    // the same rows the reference's way: set_precision through an ostringstream per value, std::endl per row
    size_t n_rows = 0;
    {
        std::ifstream in(tsv);
        std::string line;
        while (std::getline(in, line)) ++n_rows;
        n_rows -= 1;
    }
    std::vector<double> pf(n_rows), pc(n_rows);
    for (size_t i = 0; i < n_rows; ++i) {
        pf[i] = std::exp(-40.0 * rng.uniform());
        pc[i] = std::exp(-40.0 * rng.uniform());
    }
    const double t1 = wall();
    {
        std::ofstream out("/tmp/shim_bench_reference_style.tsv");
        for (size_t i = 0; i < n_rows; ++i) {
            const Snarl_data_t& sd = snarls[i];
            out << "chr1" << "\t" << sd.start_positions << "\t" << sd.end_positions << "\t"
                << std::to_string(sd.snarl_ids.first) + "_" + std::to_string(sd.snarl_ids.second) << "\t"
                << type_variants_string(sd) << "\t" << reference_set_precision(pf[i]) << "\t" << reference_set_precision(pc[i])
                << "\t" << "1234:1250,1266:1250" << "\t" << sd.depth << std::endl;
        }
    }
    const double ref_style_s = wall() - t1;
    std::ofstream js(out_json);
    js << "{\"workload\": \"" << n_snarls << " biallelic snarls x " << H << " haplotypes, binary, through GpuBinarySnarlAnalyzer\", "
       << "\"devices\": \"" << devices << "\", \"rows_written\": " << n_rows << ", \"filtered\": " << filtered
       << ", \"chunk_seconds\": " << chunk_s << ", \"stage_seconds\": {\"flatten\": " << st.flatten << ", \"set_matrix\": "
       << st.set_matrix << ", \"phenotype\": " << st.phenotype << ", \"batch\": " << st.batch << ", \"format\": " << st.format
       << ", \"write\": " << st.write << "}, \"path_tests_per_second\": " << double(2 * n_snarls) / chunk_s
       << ", \"reference_style_writer_seconds\": " << ref_style_s
       << ", \"reference_style_note\": \"the same number of rows with an ostringstream per value and std::endl per row "
          "(src/utils.cpp:5-15, src/writer.cpp:23-35), one thread\"}\n";
    std::printf("bench: %zu snarls, %zu rows, chunk %.3f s (flatten %.3f, set_matrix %.3f, batch %.3f, format %.3f, write %.3f); "
                "reference-style writer %.3f s\n", n_snarls, n_rows, chunk_s, st.flatten, st.set_matrix, st.batch, st.format,
                st.write, ref_style_s);
    return 0;
}

}  // namespace

int main(int argc, char** argv) {
    try {
        if (argc == 5 && std::strcmp(argv[1], "golden") == 0) return run_golden(argv[2], argv[3], argv[4]);
        if (argc == 6 && std::strcmp(argv[1], "bench") == 0)
            return run_bench(std::strtoull(argv[2], nullptr, 10), std::strtoull(argv[3], nullptr, 10), argv[4], argv[5]);
        if (argc == 6 && std::strcmp(argv[1], "synth") == 0)
            return run_synth(std::strtoull(argv[2], nullptr, 10), std::strtoull(argv[3], nullptr, 10),
                             std::strtoull(argv[4], nullptr, 10), argv[5]);
        std::fprintf(stderr, "usage: shim_selftest golden <golden.tsv> <pheno.tsv> <out.tsv> | synth <seed> <S> <n> <dir>\n");
        return 2;
    } catch (const std::exception& e) {  // the reference lets exceptions reach main (src/main.cpp)
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
}
