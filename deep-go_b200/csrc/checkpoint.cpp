// experiment:save / Experiment:load (experiments.lua:124-131) for this library's hosts: ONE file format for the Python and the
// Lua host.  The reference pickles the whole experiment table with torch.save; what a resumed run needs from it is the flat
// parameter vector (Torch7 getParameters order, experiments.lua:105-107) and a handful of scalars, so the format is
//     "DGCKPT02\n"  key=value lines (the library writes numLayers, channelSize, kFirst, inPlanes, P; the host appends its own:
//     rate, rateDecay, iterations, name, validation_costs ..)  "end\n"  P float32 values, little-endian.
// A checkpoint the reference itself wrote is read by dg_t7_read_model (t7_reader.cpp).
#include "model.cuh"
#include "../../include/deepgo.h"
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

extern "C" int dg_checkpoint_save(dg_model *m, const char *path, const char *meta) {
    DG_REQUIRE(m && path, "dg_checkpoint_save: NULL argument");
    std::vector<float> params(m->P);
    if (int rc = dg_model_get_params(m, params.data(), m->P)) return rc;
    if (meta)
        for (const char *q = meta; *q; ++q) DG_REQUIRE(*q != '\r', "dg_checkpoint_save: meta must be key=value lines separated by line feeds");
    const std::string tmp = std::string(path) + ".tmp";
    FILE *f = fopen(tmp.c_str(), "wb");
    DG_REQUIRE(f, "dg_checkpoint_save: cannot write %s", tmp.c_str());
    fprintf(f, "DGCKPT02\nnumLayers=%d\nchannelSize=%d\nkFirst=%d\ninPlanes=%d\nP=%zu\n", m->L, m->C, m->k_first, m->in_planes, m->P);
    if (meta && *meta) { fputs(meta, f); if (meta[strlen(meta) - 1] != '\n') fputc('\n', f); }
    fputs("end\n", f);
    const size_t w = fwrite(params.data(), 4, m->P, f);
    const int cl = fclose(f);
    DG_REQUIRE(w == m->P && cl == 0, "dg_checkpoint_save: short write to %s", tmp.c_str());
    DG_REQUIRE(rename(tmp.c_str(), path) == 0, "dg_checkpoint_save: cannot rename %s to %s", tmp.c_str(), path);   // never a half-written checkpoint under the final name
    return 0;
}

// header (everything up to and without the "end" line) into meta (NUL-terminated, at most meta_cap bytes; may be NULL), the
// parameter count into *n_out, the parameters into params when it is not NULL (cap values of room)
extern "C" int dg_checkpoint_read(const char *path, char *meta, size_t meta_cap, float *params, size_t cap, size_t *n_out) {
    DG_REQUIRE(path && n_out, "dg_checkpoint_read: NULL argument");
    FILE *f = fopen(path, "rb");
    DG_REQUIRE(f, "dg_checkpoint_read: cannot open %s", path);
    char line[4096];
    std::string hdr;
    size_t P = 0;
    bool magic = false, ended = false;
    while (fgets(line, sizeof line, f)) {
        if (!magic) {
            if (strcmp(line, "DGCKPT02\n") != 0) break;
            magic = true;
            continue;
        }
        if (strcmp(line, "end\n") == 0) { ended = true; break; }
        if (strncmp(line, "P=", 2) == 0) P = (size_t)strtoull(line + 2, nullptr, 10);
        hdr += line;
    }
    if (!magic || !ended || P == 0) { fclose(f); DG_REQUIRE(false, "dg_checkpoint_read: %s is not a DGCKPT02 checkpoint", path); }
    *n_out = P;
    if (meta && meta_cap) {
        if (hdr.size() + 1 > meta_cap) { fclose(f); DG_REQUIRE(false, "dg_checkpoint_read: header of %zu bytes does not fit %zu", hdr.size() + 1, meta_cap); }
        memcpy(meta, hdr.c_str(), hdr.size() + 1);
    }
    int rc = 0;
    if (params) {
        if (cap < P) { fclose(f); DG_REQUIRE(false, "dg_checkpoint_read: buffer holds %zu values, the file has %zu", cap, P); }
        if (fread(params, 4, P, f) != P) rc = 1;
    }
    fclose(f);
    DG_REQUIRE(rc == 0, "dg_checkpoint_read: %s is truncated", path);
    return 0;
}
