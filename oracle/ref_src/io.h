// TEST INFRASTRUCTURE: tiny binary array I/O shared by the oracle/_ref harnesses (tests/ref_io.py is the Python side).
// File = int32 count, then per array: int32 ndim, int32 dims[ndim], int32 elem_size, raw data.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <vector>
struct Arr { std::vector<int> dims; int elem = 4; std::vector<char> data; long count() const { long n = 1; for (int d : dims) n *= d; return n; }
             template <class T> T* as() { return reinterpret_cast<T*>(data.data()); } };
inline std::vector<Arr> read_arrays(const char* path) {
    FILE* f = fopen(path, "rb"); if (!f) { perror(path); exit(2); }
    int n = 0; if (fread(&n, 4, 1, f) != 1) exit(2);
    std::vector<Arr> out(n);
    for (auto& a : out) {
        int nd = 0; if (fread(&nd, 4, 1, f) != 1) exit(2);
        a.dims.resize(nd); if (nd && fread(a.dims.data(), 4, nd, f) != (size_t)nd) exit(2);
        if (fread(&a.elem, 4, 1, f) != 1) exit(2);
        a.data.resize((size_t)a.count() * a.elem);
        if (!a.data.empty() && fread(a.data.data(), 1, a.data.size(), f) != a.data.size()) exit(2);
    }
    fclose(f); return out;
}
inline void write_arrays(const char* path, const std::vector<Arr>& arrs) {
    FILE* f = fopen(path, "wb"); if (!f) { perror(path); exit(2); }
    int n = (int)arrs.size(); fwrite(&n, 4, 1, f);
    for (auto& a : arrs) {
        int nd = (int)a.dims.size(); fwrite(&nd, 4, 1, f); fwrite(a.dims.data(), 4, nd, f); fwrite(&a.elem, 4, 1, f);
        fwrite(a.data.data(), 1, a.data.size(), f);
    }
    fclose(f);
}
