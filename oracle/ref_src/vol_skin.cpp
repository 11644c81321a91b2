// TEST INFRASTRUCTURE (oracle/_ref/vol_skin): the reference's OWN trilinear skinning of the mesh by the deformed control
// lattice -- CombinedSolver::copyResultToCPUFromFloat3 / getIndex1D, examples/volumetric_mesh_deformation/src/
// CombinedSolver.h:273-276, 302-335 -- compiled from the reference's header where it lies (nothing copied).  The solver
// object is built on the CPU location with every solver switched off over a mesh of isolated vertices; its lattice, voxel
// and relative-coordinate members are then overwritten with the arrays of the test.
//   vol_skin in.bin out.bin    in: grid [(dx+1)(dy+1)(dz+1),3] f32, dims [3] i32, voxel [n,3] i32, rel [n,3] f32     out: [n,3] f32
#include <algorithm>
#include <fstream>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include "io.h"
#define private public
#define protected public
#include "CombinedSolver.h"          // -I <reference>/examples/volumetric_mesh_deformation/src
#undef private
#undef protected

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    std::vector<Arr> in = read_arrays(argv[1]);
    const int* dims = in[1].as<int>();
    const int n = in[2].dims[0];
    SimpleMesh mesh;
    for (int i = 0; i < n; ++i) mesh.add_vertex(SimpleMesh::Point((float)(i % 7), (float)(i % 5), (float)(i % 3)));   // any positions: overwritten below
    CombinedSolverParameters params;
    params.useOpt = false; params.useOptLM = false; params.useCeres = false; params.useCUDA = false;
    CombinedSolver solver(&mesh, make_int3(dims[0], dims[1], dims[2]), params, OptImage::Location::CPU, "backend_cpu", 1);
    std::vector<float3> grid((size_t)in[0].dims[0]);
    for (size_t i = 0; i < grid.size(); ++i) grid[i] = make_float3(in[0].as<float>()[3 * i], in[0].as<float>()[3 * i + 1], in[0].as<float>()[3 * i + 2]);
    solver.m_gridPosFloat3->update(grid);
    for (int i = 0; i < n; ++i) {
        solver.m_vertexToVoxels[i] = make_int3(in[2].as<int>()[3 * i], in[2].as<int>()[3 * i + 1], in[2].as<int>()[3 * i + 2]);
        solver.m_relativeCoords[i] = make_float3(in[3].as<float>()[3 * i], in[3].as<float>()[3 * i + 1], in[3].as<float>()[3 * i + 2]);
    }
    solver.copyResultToCPUFromFloat3();
    Arr out; out.dims = {n, 3}; out.elem = 4; out.data.resize((size_t)n * 12);
    int i = 0;
    for (SimpleMesh::VertexIter v = solver.m_result.vertices_begin(); v != solver.m_result.vertices_end(); ++v, ++i) {
        const SimpleMesh::Point p = solver.m_result.point(*v);
        out.as<float>()[3 * i] = p[0]; out.as<float>()[3 * i + 1] = p[1]; out.as<float>()[3 * i + 2] = p[2];
    }
    write_arrays(argv[2], {out});
    return 0;
}
