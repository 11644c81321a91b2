/* OptB200Formats.h -- dependency-free readers/writers for the on-disk formats the reference's example drivers
 * use, so the drivers' inputs can be loaded without mLib / OpenMesh / boost.
 * Host-side, plain C.  Every reader returns 0 on success, a negative code otherwise (-1 cannot open, -2
 * malformed / unsupported, -3 out of memory); buffers handed out are released with OptB200_FormatFree.
 *
 *   format                      reference that defines / consumes it
 *   .imagedump                  API/src/im.t:1-58 (writer/reader), examples/shape_from_shading/src/SimpleBuffer.cpp:17-60
 *   .SFSSolverParameters        examples/shape_from_shading/src/TerraSolverParameters.h:7-45 (raw struct: 36 floats + 3 uints)
 *   .constraints                examples/image_warping/src/main.cpp:83-111 (count, then rows "x y targetX targetY")
 *   .mrk                        examples/arap_mesh_deformation/src/main.cpp, LandMarkSet.h (count, rows "x y z radius vertex")
 *   .ply / .off meshes          read by the mesh drivers through OpenMesh (triangle meshes: vertex positions + faces)
 */
#ifndef OPT_B200_FORMATS_H
#define OPT_B200_FORMATS_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int width, height, channels;
    int datatype;              /* 0: float32, 1: uint8 (SimpleBuffer.h:7) */
    void* data;                /* row-major, width fastest, no padding */
} OptB200_ImageDump;

/* clamp_infinity != 0 reproduces SimpleBuffer's default (+inf -> FLT_MAX, -inf -> -10000.0f over the first
 * width*height floats, SimpleBuffer.cpp:34-46). */
int  OptB200_ImageDumpRead(const char* path, int clamp_infinity, OptB200_ImageDump* out);
int  OptB200_ImageDumpWrite(const char* path, const OptB200_ImageDump* image);

typedef struct {               /* TerraSolverParameters.h:7-31, field for field */
    float weightFitting, weightRegularizer, weightPrior, weightShading, weightShadingStart, weightShadingIncrement,
          weightBoundary, fx, fy, ux, uy;
    float deltaTransform[16];
    float lightingCoefficients[9];
    unsigned int unused[3];
} OptB200_SFSParameters;
int  OptB200_SFSParametersRead(const char* path, OptB200_SFSParameters* out);

/* *rows receives count x 4 ints (x, y, targetX, targetY). */
int  OptB200_ConstraintsRead(const char* path, int* count, int** rows);
/* *positions receives count x 3 floats, *vertices count ints (the radius column is dropped). */
int  OptB200_MarkersRead(const char* path, int* count, float** positions, int** vertices);

/* Triangle meshes: PLY (ascii or binary_little_endian; float/double vertex properties x,y,z anywhere in the
 * vertex element, list-typed faces with 3 indices) and OFF (ascii).  *vertices: nv x 3 floats, *faces: nf x 3 ints. */
int  OptB200_MeshRead(const char* path, int* nv, float** vertices, int* nf, int** faces);
/* ascii OFF (the format the mesh drivers write their results in, e.g. out.off). */
int  OptB200_MeshWriteOFF(const char* path, int nv, const float* vertices, int nf, const int* faces);

void OptB200_FormatFree(void* p);

#ifdef __cplusplus
}
#endif
#endif
