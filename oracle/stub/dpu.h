/*
 * TEST INFRASTRUCTURE ONLY -- stand-in for the UPMEM host SDK header <dpu.h>.
 *
 * The reference's snappy_compress.c / snappy_decompress.c keep the host codec
 * (the oracle) and the DPU orchestration in one translation unit and include
 * <dpu.h>, <dpu_memory.h>, <dpu_log.h> (snappy_compress.c:1-3,
 * snappy_decompress.c:1-3).  The SDK is not installed here, so to compile the
 * host codec from the reference sources *where they lie* this header supplies
 * just enough declarations for the DPU half to compile.  Every SDK entry point
 * is a no-op and dpu_launch() reports failure, so the DPU path can never
 * appear to work.  Nothing in the shipped product includes this file.
 */
#ifndef ORACLE_STUB_DPU_H
#define ORACLE_STUB_DPU_H

#include <stdbool.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

struct dpu_set_t { int unused; };
typedef int dpu_error_t;

#define DPU_OK 0
#define DPU_XFER_TO_DPU 0
#define DPU_XFER_FROM_DPU 1
#define DPU_XFER_DEFAULT 0
#define DPU_SYNCHRONOUS 0

#define DPU_ASSERT(stmt) do { (void)(stmt); } while (0)
/* zero-trip loops: there are no ranks and no DPUs */
#define DPU_RANK_FOREACH(set, rank) for ((rank) = (set); 0;)
#define DPU_FOREACH(set, dpu) for ((dpu) = (set); 0;)

static inline dpu_error_t dpu_alloc(uint32_t n, const char *p, struct dpu_set_t *s) { (void)n; (void)p; s->unused = 0; return DPU_OK; }
static inline dpu_error_t dpu_load(struct dpu_set_t s, const char *f, void *p) { (void)s; (void)f; (void)p; return DPU_OK; }
static inline dpu_error_t dpu_prepare_xfer(struct dpu_set_t s, void *b) { (void)s; (void)b; return DPU_OK; }
static inline dpu_error_t dpu_push_xfer(struct dpu_set_t s, int dir, const char *sym, uint32_t off, size_t len, int fl) { (void)s; (void)dir; (void)sym; (void)off; (void)len; (void)fl; return DPU_OK; }
static inline dpu_error_t dpu_copy_to(struct dpu_set_t s, const char *sym, uint32_t off, const void *src, size_t len) { (void)s; (void)sym; (void)off; (void)src; (void)len; return DPU_OK; }
static inline dpu_error_t dpu_copy_from(struct dpu_set_t s, const char *sym, uint32_t off, void *dst, size_t len) { (void)s; (void)sym; (void)off; (void)dst; (void)len; return DPU_OK; }
static inline dpu_error_t dpu_launch(struct dpu_set_t s, int pol) { (void)s; (void)pol; return 1; /* always fails */ }
static inline dpu_error_t dpu_free(struct dpu_set_t s) { (void)s; return DPU_OK; }
static inline dpu_error_t dpu_get_nr_dpus(struct dpu_set_t s, uint32_t *n) { (void)s; *n = 0; return DPU_OK; }
static inline dpu_error_t dpu_log_read(struct dpu_set_t s, FILE *f) { (void)s; (void)f; return DPU_OK; }

#endif
