/* formats.c -- readers/writers for the reference drivers' on-disk formats (include/OptB200Formats.h). */
#include <ctype.h>
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/OptB200Formats.h"

enum { E_OPEN = -1, E_FORMAT = -2, E_MEMORY = -3 };

void OptB200_FormatFree(void* p) { free(p); }

/* ---- .imagedump: int32 width, height, channelCount, datatype, then the pixels (im.t:9-15) ---------------- */
int OptB200_ImageDumpRead(const char* path, int clamp_infinity, OptB200_ImageDump* out) {
    FILE* f = fopen(path, "rb");
    if (!f) return E_OPEN;
    int32_t hdr[4];
    if (fread(hdr, sizeof(int32_t), 4, f) != 4 || hdr[0] <= 0 || hdr[1] <= 0 || hdr[2] <= 0 || (hdr[3] != 0 && hdr[3] != 1)) {
        fclose(f);
        return E_FORMAT;
    }
    const size_t elem = hdr[3] == 0 ? sizeof(float) : 1;
    /* three untrusted 31-bit factors: their product overflows size_t.  Bound it by what the file can hold before malloc. */
    long here = ftell(f), end = -1;
    if (here >= 0 && fseek(f, 0, SEEK_END) == 0) { end = ftell(f); if (fseek(f, here, SEEK_SET) != 0) end = -1; }
    if (here < 0 || end < here) { fclose(f); return E_FORMAT; }
    const unsigned long long avail = (unsigned long long)(end - here) / elem;
    const unsigned long long wh = (unsigned long long)hdr[0] * (unsigned long long)hdr[1];        /* < 2^62 */
    if (wh > avail || wh * (unsigned long long)hdr[2] > avail) { fclose(f); return E_FORMAT; }
    const size_t count = (size_t)(wh * (unsigned long long)hdr[2]);
    void* data = malloc(count * elem);
    if (!data) { fclose(f); return E_MEMORY; }
    if (fread(data, elem, count, f) != count) { free(data); fclose(f); return E_FORMAT; }
    fclose(f);
    if (hdr[3] == 0 && clamp_infinity) {
        float* p = (float*)data;
        const size_t n = (size_t)hdr[0] * (size_t)hdr[1];     /* SimpleBuffer.cpp:36 walks width*height floats */
        for (size_t i = 0; i < n; ++i)
            if (isinf(p[i])) p[i] = p[i] > 0 ? FLT_MAX : -10000.0f;
    }
    out->width = hdr[0]; out->height = hdr[1]; out->channels = hdr[2]; out->datatype = hdr[3]; out->data = data;
    return 0;
}

int OptB200_ImageDumpWrite(const char* path, const OptB200_ImageDump* im) {
    if (!im || !im->data || (im->datatype != 0 && im->datatype != 1)) return E_FORMAT;
    FILE* f = fopen(path, "wb");
    if (!f) return E_OPEN;
    const int32_t hdr[4] = {im->width, im->height, im->channels, im->datatype};
    const size_t elem = im->datatype == 0 ? sizeof(float) : 1;
    const size_t count = (size_t)im->width * (size_t)im->height * (size_t)im->channels;
    int ok = fwrite(hdr, sizeof(int32_t), 4, f) == 4 && fwrite(im->data, elem, count, f) == count;
    ok = (fclose(f) == 0) && ok;
    return ok ? 0 : E_FORMAT;
}

int OptB200_SFSParametersRead(const char* path, OptB200_SFSParameters* out) {
    FILE* f = fopen(path, "rb");
    if (!f) return E_OPEN;
    const size_t need = 36 * sizeof(float) + 3 * sizeof(unsigned int);   /* files carry 4 bytes of struct padding after */
    unsigned char buf[36 * sizeof(float) + 3 * sizeof(unsigned int)];
    const size_t got = fread(buf, 1, need, f);
    fclose(f);
    if (got != need) return E_FORMAT;
    memcpy(out, buf, need);
    return 0;
}

/* ---- whitespace-separated text tables ----------------------------------------------------------------------- */
static char* slurp(const char* path, size_t* len) {
    FILE* f = fopen(path, "rb");
    if (!f) return NULL;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    char* s = n >= 0 ? (char*)malloc((size_t)n + 1) : NULL;
    if (s && fread(s, 1, (size_t)n, f) != (size_t)n) { free(s); s = NULL; }
    fclose(f);
    if (s) { s[n] = 0; if (len) *len = (size_t)n; }
    return s;
}
static int next_double(char** cur, double* v) {
    char* end;
    *v = strtod(*cur, &end);
    if (end == *cur) return 0;
    *cur = end;
    return 1;
}

int OptB200_ConstraintsRead(const char* path, int* count, int** rows) {
    char* txt = slurp(path, NULL);
    if (!txt) return E_OPEN;
    char* cur = txt;
    double v;
    if (!next_double(&cur, &v) || v < 0) { free(txt); return E_FORMAT; }
    const int n = (int)v;
    int* r = (int*)malloc(sizeof(int) * 4 * (size_t)(n > 0 ? n : 1));
    if (!r) { free(txt); return E_MEMORY; }
    for (int i = 0; i < 4 * n; ++i) {
        if (!next_double(&cur, &v)) { free(r); free(txt); return E_FORMAT; }
        r[i] = (int)v;
    }
    free(txt);
    *count = n; *rows = r;
    return 0;
}

int OptB200_MarkersRead(const char* path, int* count, float** positions, int** vertices) {
    char* txt = slurp(path, NULL);
    if (!txt) return E_OPEN;
    char* cur = txt;
    double v;
    if (!next_double(&cur, &v) || v < 0) { free(txt); return E_FORMAT; }
    const int n = (int)v;
    float* p = (float*)malloc(sizeof(float) * 3 * (size_t)(n > 0 ? n : 1));
    int* idx = (int*)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    if (!p || !idx) { free(p); free(idx); free(txt); return E_MEMORY; }
    for (int i = 0; i < n; ++i) {
        double row[5];
        for (int k = 0; k < 5; ++k)
            if (!next_double(&cur, &row[k])) { free(p); free(idx); free(txt); return E_FORMAT; }
        p[3 * i] = (float)row[0]; p[3 * i + 1] = (float)row[1]; p[3 * i + 2] = (float)row[2];
        idx[i] = (int)row[4];
    }
    free(txt);
    *count = n; *positions = p; *vertices = idx;
    return 0;
}

/* ---- meshes ----------------------------------------------------------------------------------------------- */
static int type_size(const char* t) {
    if (!strcmp(t, "char") || !strcmp(t, "uchar") || !strcmp(t, "int8") || !strcmp(t, "uint8")) return 1;
    if (!strcmp(t, "short") || !strcmp(t, "ushort") || !strcmp(t, "int16") || !strcmp(t, "uint16")) return 2;
    if (!strcmp(t, "int") || !strcmp(t, "uint") || !strcmp(t, "float") || !strcmp(t, "int32") || !strcmp(t, "uint32") || !strcmp(t, "float32")) return 4;
    if (!strcmp(t, "double") || !strcmp(t, "float64")) return 8;
    return 0;
}
static int is_float_type(const char* t) { return !strcmp(t, "float") || !strcmp(t, "float32") || !strcmp(t, "double") || !strcmp(t, "float64"); }
static double read_scalar(const unsigned char* p, const char* t) {
    switch (type_size(t)) {
    case 1: return t[0] == 'u' ? (double)*(const uint8_t*)p : (double)*(const int8_t*)p;
    case 2: { if (t[0] == 'u') { uint16_t v; memcpy(&v, p, 2); return v; } int16_t v; memcpy(&v, p, 2); return v; }
    case 4: if (is_float_type(t)) { float v; memcpy(&v, p, 4); return v; }
            if (t[0] == 'u') { uint32_t v; memcpy(&v, p, 4); return v; } else { int32_t v; memcpy(&v, p, 4); return v; }
    case 8: { double v; memcpy(&v, p, 8); return v; }
    }
    return 0;
}

#define MAX_PROPS 32
typedef struct { char type[16]; char name[32]; int is_list; char count_type[16]; } PlyProp;

static int read_ply(char* txt, size_t len, int* nv_out, float** v_out, int* nf_out, int** f_out) {
    int binary = 0, nv = 0, nf = 0, nvp = 0, nfp = 0, section = 0;     /* section: 1 vertex, 2 face, 3 other */
    PlyProp vp[MAX_PROPS], fp[MAX_PROPS];
    char* cur = txt;
    char* body = NULL;
    while (cur < txt + len) {
        char* eol = memchr(cur, '\n', (size_t)(txt + len - cur));
        if (!eol) return E_FORMAT;
        size_t n = (size_t)(eol - cur);
        char line[256];
        if (n >= sizeof(line)) n = sizeof(line) - 1;
        memcpy(line, cur, n);
        line[n] = 0;
        while (n && isspace((unsigned char)line[n - 1])) line[--n] = 0;
        cur = eol + 1;
        char a[32] = "", b[32] = "", c[32] = "", d[32] = "";
        int k = sscanf(line, "%31s %31s %31s %31s", a, b, c, d);
        if (k < 1) continue;
        if (!strcmp(a, "format")) {
            if (!strcmp(b, "ascii")) binary = 0;
            else if (!strcmp(b, "binary_little_endian")) binary = 1;
            else return E_FORMAT;
        } else if (!strcmp(a, "element")) {
            if (!strcmp(b, "vertex")) { nv = atoi(c); section = 1; }
            else if (!strcmp(b, "face")) { nf = atoi(c); section = 2; }
            else { if (atoi(c) != 0 && (section == 0 || nf == 0)) return E_FORMAT; section = 3; }   /* elements before the faces must be empty */
        } else if (!strcmp(a, "property")) {
            PlyProp pr;
            memset(&pr, 0, sizeof(pr));
            if (!strcmp(b, "list")) {
                char e[32] = "";
                sscanf(line, "%*s %*s %15s %15s %31s", pr.count_type, pr.type, e);
                pr.is_list = 1;
                strncpy(pr.name, e, sizeof(pr.name) - 1);
            } else {
                strncpy(pr.type, b, sizeof(pr.type) - 1);
                strncpy(pr.name, c, sizeof(pr.name) - 1);
            }
            if (section == 1 && nvp < MAX_PROPS) vp[nvp++] = pr;
            else if (section == 2 && nfp < MAX_PROPS) fp[nfp++] = pr;
        } else if (!strcmp(a, "end_header")) { body = cur; break; }
    }
    if (!body || nv <= 0 || nf < 0) return E_FORMAT;
    float* V = (float*)malloc(sizeof(float) * 3 * (size_t)nv);
    int* F = (int*)malloc(sizeof(int) * 3 * (size_t)(nf > 0 ? nf : 1));
    if (!V || !F) { free(V); free(F); return E_MEMORY; }
    int rc = 0;
    if (binary) {
        const unsigned char* p = (const unsigned char*)body;
        const unsigned char* end = (const unsigned char*)txt + len;
        for (int i = 0; i < nv && !rc; ++i)
            for (int j = 0; j < nvp; ++j) {
                const int sz = type_size(vp[j].type);
                if (vp[j].is_list || !sz || p + sz > end) { rc = E_FORMAT; break; }
                const double val = read_scalar(p, vp[j].type);
                if (!strcmp(vp[j].name, "x")) V[3 * i] = (float)val;
                else if (!strcmp(vp[j].name, "y")) V[3 * i + 1] = (float)val;
                else if (!strcmp(vp[j].name, "z")) V[3 * i + 2] = (float)val;
                p += sz;
            }
        for (int i = 0; i < nf && !rc; ++i)
            for (int j = 0; j < nfp; ++j) {
                if (!fp[j].is_list) { const int sz = type_size(fp[j].type); if (!sz || p + sz > end) { rc = E_FORMAT; break; } p += sz; continue; }
                const int cs = type_size(fp[j].count_type), es = type_size(fp[j].type);
                if (!cs || !es || p + cs > end) { rc = E_FORMAT; break; }
                const int cnt = (int)read_scalar(p, fp[j].count_type);
                p += cs;
                if (cnt < 0 || p + (size_t)cnt * es > end) { rc = E_FORMAT; break; }
                if (j == 0) {
                    if (cnt != 3) { rc = E_FORMAT; break; }
                    for (int q = 0; q < 3; ++q) F[3 * i + q] = (int)read_scalar(p + q * es, fp[j].type);
                }
                p += (size_t)cnt * es;
            }
    } else {
        char* c2 = body;
        double val;
        for (int i = 0; i < nv && !rc; ++i)
            for (int j = 0; j < nvp; ++j) {
                if (!next_double(&c2, &val)) { rc = E_FORMAT; break; }
                if (!strcmp(vp[j].name, "x")) V[3 * i] = (float)val;
                else if (!strcmp(vp[j].name, "y")) V[3 * i + 1] = (float)val;
                else if (!strcmp(vp[j].name, "z")) V[3 * i + 2] = (float)val;
            }
        for (int i = 0; i < nf && !rc; ++i) {
            if (!next_double(&c2, &val) || (int)val != 3) { rc = E_FORMAT; break; }
            for (int q = 0; q < 3; ++q) { if (!next_double(&c2, &val)) { rc = E_FORMAT; break; } F[3 * i + q] = (int)val; }
            char* eol = strchr(c2, '\n');                     /* further per-face properties on the same line */
            if (eol) c2 = eol;
        }
    }
    if (rc) { free(V); free(F); return rc; }
    *nv_out = nv; *v_out = V; *nf_out = nf; *f_out = F;
    return 0;
}

static int read_off(char* txt, int* nv_out, float** v_out, int* nf_out, int** f_out) {
    /* drop comments */
    for (char* c = txt; *c; ++c)
        if (*c == '#') {
            while (*c && *c != '\n') *c++ = ' ';
            if (!*c) break;                                   /* comment ran to the end of the file: do not step past the NUL */
        }
    char* cur = txt;
    while (isspace((unsigned char)*cur)) ++cur;
    if (strncmp(cur, "OFF", 3) != 0) return E_FORMAT;
    cur += 3;
    double a, b, e;
    if (!next_double(&cur, &a) || !next_double(&cur, &b) || !next_double(&cur, &e) || a <= 0 || b < 0) return E_FORMAT;
    const int nv = (int)a, nf = (int)b;
    float* V = (float*)malloc(sizeof(float) * 3 * (size_t)nv);
    int* F = (int*)malloc(sizeof(int) * 3 * (size_t)(nf > 0 ? nf : 1));
    if (!V || !F) { free(V); free(F); return E_MEMORY; }
    double val;
    for (int i = 0; i < 3 * nv; ++i) { if (!next_double(&cur, &val)) { free(V); free(F); return E_FORMAT; } V[i] = (float)val; }
    for (int i = 0; i < nf; ++i) {
        if (!next_double(&cur, &val) || (int)val != 3) { free(V); free(F); return E_FORMAT; }
        for (int q = 0; q < 3; ++q) { if (!next_double(&cur, &val)) { free(V); free(F); return E_FORMAT; } F[3 * i + q] = (int)val; }
        char* eol = strchr(cur, '\n');                        /* optional colour columns */
        if (eol) cur = eol;
    }
    *nv_out = nv; *v_out = V; *nf_out = nf; *f_out = F;
    return 0;
}

int OptB200_MeshRead(const char* path, int* nv, float** vertices, int* nf, int** faces) {
    size_t len = 0;
    char* txt = slurp(path, &len);
    if (!txt) return E_OPEN;
    int rc;
    if (len >= 3 && !strncmp(txt, "ply", 3)) rc = read_ply(txt, len, nv, vertices, nf, faces);
    else rc = read_off(txt, nv, vertices, nf, faces);
    free(txt);
    return rc;
}

int OptB200_MeshWriteOFF(const char* path, int nv, const float* v, int nf, const int* f) {
    FILE* o = fopen(path, "w");
    if (!o) return E_OPEN;
    fprintf(o, "OFF\n%d %d 0\n", nv, nf);
    for (int i = 0; i < nv; ++i) fprintf(o, "%.9g %.9g %.9g\n", v[3 * i], v[3 * i + 1], v[3 * i + 2]);
    for (int i = 0; i < nf; ++i) fprintf(o, "3 %d %d %d\n", f[3 * i], f[3 * i + 1], f[3 * i + 2]);
    return fclose(o) == 0 ? 0 : E_FORMAT;
}
