/*
 * pim_synth.h -- seeded, integer-only workload generators (libpim_synth.so), bit-identical on host and
 * device; definitions in DESIGN.md and csrc/synth.cu.  They stand in for the corpus files BASELINE.json
 * names that are absent from the reference checkout.  Benchmark / test support: not part
 * of the codec library, so that a process can generate inputs without loading libpim_snappy.so.
 *
 * kind: 0 text-like, 1 random, 2 mixed (1 MiB text / 1 MiB random alternating), 3 binary-like mix.
 */
#ifndef PIM_SYNTH_H
#define PIM_SYNTH_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { PIM_SYNTH_TEXT = 0, PIM_SYNTH_RANDOM = 1, PIM_SYNTH_MIX = 2, PIM_SYNTH_BIN = 3 };
/* 0 on success */
int pim_snappy_synth_host(int kind, uint64_t seed, uint8_t *out, uint64_t n);
int pim_snappy_synth_device(int kind, uint64_t seed, uint8_t *d_out, uint64_t n, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PIM_SYNTH_H */
