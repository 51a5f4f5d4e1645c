// preset2bin — dump the flat EsbEngineParams blob (and IR file names) of a preset.
// Used by tests/golden/make_golden.py to derive fixtures that travel to the GPU box.
#include "host.h"
#include <cstdio>
int main(int argc, char **argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: preset2bin <preset.json> <out.bin>\n"); return 2; }
    esb::LoadedPreset lp;
    try { esb::loadEngineFile(argv[1], lp); } catch (const std::exception &e) { std::fprintf(stderr, "error: %s\n", e.what()); return 1; }
    FILE *f = std::fopen(argv[2], "wb");
    if (!f) return 1;
    std::fwrite(&lp.params, sizeof(lp.params), 1, f);
    std::fclose(f);
    std::printf("%s cylinders=%d functions=%d bytes=%zu\n", lp.name.c_str(), lp.params.cylinderCount, lp.params.functionCount, sizeof(lp.params));
    for (auto &s : lp.irFiles) std::printf("ir %s\n", s.c_str());
    return 0;
}
