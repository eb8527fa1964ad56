// grb_storage.cuh — storage plumbing shared between translation units (see grb_storage.cu)
#pragma once
#include "grb_internal.cuh"

Csr coo_to_csr(int64_t nrows, int64_t ncols, uint64_t *keys, const void *vals, bool iso_in, int tcode,
               int64_t n, int dup_op, bool keep_iso);
uint32_t *csr_row_ids(const Csr &c);                       // expanded row index per entry (device, caller frees)
void narrow_indices(const uint64_t *d_in, int64_t n, uint32_t *d_out, uint64_t bound, const char *what);
void widen_indices(const uint32_t *d_in, int64_t n, uint64_t *d_out);
bool csr_is_sorted(const Csr &c);                          // also validates rowptr; throws on corrupt rowptr
void csr_sort_rows(Csr &c, int tcode);                     // sort columns within rows (drops nothing)
int64_t count_present(const uint8_t *present, int64_t n, int64_t *d_count);   // enqueues; returns nothing useful
void cast_array(const void *in, int from_code, void *out, int to_code, int64_t n);   // C typecast rules
