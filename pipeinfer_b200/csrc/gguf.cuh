// gguf.cuh -- minimal GGUF v2/v3 reader (host).  Replaces, for the path only, the reference's llama_model_loader
// (llama.cpp:1805-2138) + gguf parser (ggml.c:18083-18300): header, key/value section, tensor infos, aligned data
// section; the file is memory-mapped and tensors are uploaded straight from the mapping.
#pragma once

#include <fcntl.h>
#include <stdint.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <map>
#include <string>
#include <vector>

struct GgufTensor {
    std::string name;
    int         n_dims = 0;
    int64_t     ne[4]  = { 1, 1, 1, 1 };
    int         type   = 0;
    uint64_t    offset = 0;
    const void * data  = nullptr;
};

struct GgufValue {
    int                      type = -1;       // gguf value type
    uint64_t                 u    = 0;
    int64_t                  i    = 0;
    double                   f    = 0;
    std::string              s;
    int                      arr_type = -1;
    std::vector<std::string> arr_s;
    std::vector<double>      arr_f;
    std::vector<int64_t>     arr_i;
};

struct GgufFile {
    int                               fd   = -1;
    uint8_t *                         base = nullptr;
    size_t                            size = 0;
    uint32_t                          version = 0;
    std::map<std::string, GgufValue>  kv;
    std::map<std::string, GgufTensor> tensors;
    std::string                       error;

    ~GgufFile() {
        if (base) munmap(base, size);
        if (fd >= 0) close(fd);
    }

    bool open(const char * path) {
        fd = ::open(path, O_RDONLY);
        if (fd < 0) { error = std::string("cannot open ") + path; return false; }
        struct stat sb;
        fstat(fd, &sb);
        size = (size_t) sb.st_size;
        base = (uint8_t *) mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (base == MAP_FAILED) { base = nullptr; error = "mmap failed"; return false; }
        size_t p = 0;
        auto   need = [&](size_t n) { return p + n <= size; };
        auto   rd32 = [&]() { uint32_t v; memcpy(&v, base + p, 4); p += 4; return v; };
        auto   rd64 = [&]() { uint64_t v; memcpy(&v, base + p, 8); p += 8; return v; };
        auto   rdstr = [&]() { uint64_t n = rd64(); std::string s((const char *) base + p, (size_t) n); p += n; return s; };
        if (!need(24) || memcmp(base, "GGUF", 4) != 0) { error = "not a GGUF file"; return false; }
        p       = 4;
        version = rd32();
        if (version < 2 || version > 3) { error = "unsupported GGUF version"; return false; }
        const uint64_t n_tensors = rd64(), n_kv = rd64();
        static const int scalar_size[] = { 1, 1, 2, 2, 4, 4, 4, 1, 0, 0, 8, 8, 8 };
        auto rdscalar = [&](int t, GgufValue & v) {
            switch (t) {
                case 0: v.u = base[p]; v.i = (int64_t) v.u; break;
                case 1: v.i = (int8_t) base[p]; v.u = (uint64_t) v.i; break;
                case 2: { uint16_t x; memcpy(&x, base + p, 2); v.u = x; v.i = x; } break;
                case 3: { int16_t x; memcpy(&x, base + p, 2); v.i = x; v.u = (uint64_t) x; } break;
                case 4: { uint32_t x; memcpy(&x, base + p, 4); v.u = x; v.i = x; } break;
                case 5: { int32_t x; memcpy(&x, base + p, 4); v.i = x; v.u = (uint64_t) x; } break;
                case 6: { float x; memcpy(&x, base + p, 4); v.f = x; } break;
                case 7: v.u = base[p] != 0; v.i = (int64_t) v.u; break;
                case 10: { uint64_t x; memcpy(&x, base + p, 8); v.u = x; v.i = (int64_t) x; } break;
                case 11: { int64_t x; memcpy(&x, base + p, 8); v.i = x; v.u = (uint64_t) x; } break;
                case 12: { double x; memcpy(&x, base + p, 8); v.f = x; } break;
            }
            if (t != 6 && t != 12) v.f = (double) v.i;
            p += scalar_size[t];
        };
        for (uint64_t k = 0; k < n_kv; k++) {
            std::string key = rdstr();
            GgufValue   v;
            v.type = (int) rd32();
            if (v.type == 8) {
                v.s = rdstr();
            } else if (v.type == 9) {
                v.arr_type       = (int) rd32();
                const uint64_t n = rd64();
                for (uint64_t e = 0; e < n; e++) {
                    if (v.arr_type == 8) v.arr_s.push_back(rdstr());
                    else if (v.arr_type <= 12 && v.arr_type != 9) {
                        GgufValue t;
                        rdscalar(v.arr_type, t);
                        if (v.arr_type == 6 || v.arr_type == 12) v.arr_f.push_back(t.f);
                        else v.arr_i.push_back(t.i);
                    } else { error = "unsupported array type"; return false; }
                }
            } else if (v.type >= 0 && v.type <= 12) {
                rdscalar(v.type, v);
            } else { error = "bad value type"; return false; }
            kv[key] = v;
        }
        std::vector<GgufTensor> infos;
        for (uint64_t t = 0; t < n_tensors; t++) {
            GgufTensor ti;
            ti.name   = rdstr();
            ti.n_dims = (int) rd32();
            for (int d = 0; d < ti.n_dims; d++) ti.ne[d] = (int64_t) rd64();
            ti.type   = (int) rd32();
            ti.offset = rd64();
            infos.push_back(ti);
        }
        uint64_t align = 32;
        if (kv.count("general.alignment")) align = kv["general.alignment"].u;
        const size_t data_off = (p + align - 1) / align * align;
        for (auto & ti : infos) {
            if (data_off + ti.offset > size) { error = "tensor offset out of range: " + ti.name; return false; }
            ti.data          = base + data_off + ti.offset;
            tensors[ti.name] = ti;
        }
        return true;
    }

    int64_t     get_i(const std::string & k, int64_t def) const { auto it = kv.find(k); return it == kv.end() ? def : it->second.i; }
    double      get_f(const std::string & k, double def) const { auto it = kv.find(k); return it == kv.end() ? def : it->second.f; }
    std::string get_s(const std::string & k, const std::string & def) const { auto it = kv.find(k); return it == kv.end() ? def : it->second.s; }
    const GgufTensor * tensor(const std::string & name) const { auto it = tensors.find(name); return it == tensors.end() ? nullptr : &it->second; }
};
