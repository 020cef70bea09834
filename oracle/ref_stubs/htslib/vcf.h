/* htslib stand-in.  TEST INFRASTRUCTURE (see ../README.md).  Every function aborts: nothing on the path calls them. */
#pragma once
#include "hts.h"
struct bcf_hdr_t { int unused; };
struct bcf1_t { int32_t rid; int64_t pos; int n_allele; uint32_t n_sample; struct { char** allele; char* id; } d; };
#define BCF_UN_STR 1
#define BCF_UN_ALL 15
#define BCF_DT_SAMPLE 2
#define BCF_DT_ID 0
#define BCF_DT_CTG 1
#define bcf_gt_allele(val) (((val) >> 1) - 1)
#define bcf_gt_is_missing(val) ((val) >> 1 ? 0 : 1)
#define bcf_gt_is_phased(val) ((val) & 1)
#define bcf_int32_vector_end (-2147483647)
#define bcf_int32_missing (-2147483647 - 1)
[[noreturn]] inline void stoat_ref_stub_hts() { abort(); }
inline htsFile* bcf_open(const char*, const char*) { stoat_ref_stub_hts(); }
inline int bcf_close(htsFile*) { stoat_ref_stub_hts(); }
inline bcf_hdr_t* bcf_hdr_read(htsFile*) { stoat_ref_stub_hts(); }
inline void bcf_hdr_destroy(bcf_hdr_t*) { stoat_ref_stub_hts(); }
inline bcf1_t* bcf_init() { stoat_ref_stub_hts(); }
inline void bcf_destroy(bcf1_t*) { stoat_ref_stub_hts(); }
inline int bcf_read(htsFile*, const bcf_hdr_t*, bcf1_t*) { stoat_ref_stub_hts(); }
inline int bcf_unpack(bcf1_t*, int) { stoat_ref_stub_hts(); }
inline const char* bcf_hdr_id2name(const bcf_hdr_t*, int) { stoat_ref_stub_hts(); }
inline const char* bcf_hdr_int2id(const bcf_hdr_t*, int, int) { stoat_ref_stub_hts(); }
inline int bcf_hdr_nsamples(const bcf_hdr_t*) { stoat_ref_stub_hts(); }
inline int bcf_get_info_int32(const bcf_hdr_t*, bcf1_t*, const char*, int32_t**, int*) { stoat_ref_stub_hts(); }
inline int bcf_get_info_string(const bcf_hdr_t*, bcf1_t*, const char*, char**, int*) { stoat_ref_stub_hts(); }
inline int bcf_get_genotypes(const bcf_hdr_t*, bcf1_t*, int32_t**, int*) { stoat_ref_stub_hts(); }
inline int bcf_get_format_int32(const bcf_hdr_t*, bcf1_t*, const char*, int32_t**, int*) { stoat_ref_stub_hts(); }
