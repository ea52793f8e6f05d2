// corpus_host.cpp — TEST / BENCH INFRASTRUCTURE.  The synthetic corpus generator of BASELINE.json's
// configs 2, 4 and 5 (shellac_b200/csrc/corpus.cuh, a counter-based generator written to compile
// for host and device alike) built as a plain host library, so that the CPU arms of bench.py
// (`cpu_baseline`, `--impl reference`) and the CPU tests get byte-identical inputs WITHOUT loading
// the product library libshellac_b200.so.  Nothing of the hot path is in here.
#include <cstdint>
#include "../shellac_b200/csrc/corpus.cuh"

extern "C" {
int orc_corpus_lengths(uint64_t seed, uint64_t first, uint64_t n, int kind, uint32_t* len) {
    if (!len) return -1;
    for (uint64_t i = 0; i < n; i++) len[i] = shl::gen_message(seed, first + i, kind, nullptr);
    return 0;
}
int orc_corpus_fill(uint64_t seed, uint64_t first, uint64_t n, int kind, const uint64_t* off, uint8_t* arena) {
    if (!off || !arena) return -1;
    for (uint64_t i = 0; i < n; i++) shl::gen_message(seed, first + i, kind, arena + off[i]);
    return 0;
}
}
