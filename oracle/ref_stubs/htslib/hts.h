/* htslib stand-in: opaque types and never-called functions.  TEST INFRASTRUCTURE (see ../README.md).
 * VCF reading (make_edge_matrix, src/snarl_analyzer.cpp:190-260) is outside the path; the harness fills the matrix. */
#pragma once
#include <cstdint>
#include <cstdlib>
struct htsFile { int unused; };
struct hts_idx_t;
