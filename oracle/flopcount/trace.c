/* trace.c — basic-block execution counter behind -fsanitize-coverage=trace-pc (TEST INFRASTRUCTURE).
 * Every instrumented basic block calls __sanitizer_cov_trace_pc(); while counting is enabled the
 * return address is tallied in an open-addressing table, dumped as "pc count" lines. */
#include <stdint.h>
#include <stdio.h>
#define TABLE (1u << 20)
static uintptr_t keys[TABLE];
static uint64_t counts[TABLE];
int flopcount_enabled = 0;
void __sanitizer_cov_trace_pc(void) {
    if (!flopcount_enabled) return;
    const uintptr_t pc = (uintptr_t)__builtin_return_address(0);
    uint32_t h = (uint32_t)((pc * 0x9E3779B97F4A7C15ull) >> 44) & (TABLE - 1);
    while (keys[h] != 0 && keys[h] != pc) h = (h + 1) & (TABLE - 1);
    keys[h] = pc;
    counts[h]++;
}
void flopcount_dump(const char *path) {
    FILE *f = fopen(path, "w");
    for (uint32_t i = 0; i < TABLE; ++i)
        if (keys[i]) fprintf(f, "%lx %llu\n", (unsigned long)keys[i], (unsigned long long)counts[i]);
    fclose(f);
}
