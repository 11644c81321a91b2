// TEST INFRASTRUCTURE.  A caller written against the reference's OWN driver layer -- the unmodified headers
// examples/shared/{OptSolver.h, OptUtils.h, NamedParameters.h, OptImage.h, OptGraph.h, CombinedSolverParameters.h},
// included from /root/reference where they lie -- with include/Opt.h first on the include path and libOptB200.so on
// the link line (the one-line change of examples/shared/make_template.inc).  It reads a problem dump
// written by tests/test_ref_driver.py, builds OptImage / OptGraph / NamedParameters exactly as the examples'
// CombinedSolver classes do, runs OptSolver::solve (launchProfiledSolve: Opt_ProblemInit + Opt_ProblemStep loop) and
// prints the final cost; the library itself prints `final cost=` as the reference does (verbosityLevel 1).
//
//   ref_harness <problem.bin> [--materialized] [--fused] [--lm]
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "OptSolver.h"          // the reference's: pulls in Opt.h (ours, first on the path), OptUtils.h, NamedParameters.h ...
#include "CombinedSolverParameters.h"

static unsigned rd_u32(FILE* f) { unsigned v = 0; if (fread(&v, 4, 1, f) != 1) { fprintf(stderr, "short read\n"); exit(2); } return v; }
static std::string rd_str(FILE* f) { unsigned n = rd_u32(f); std::string s(n, '\0'); if (n && fread(&s[0], 1, n, f) != n) exit(2); return s; }

int main(int argc, const char* argv[]) {
    if (argc < 2) { fprintf(stderr, "usage: ref_harness problem.bin [--materialized] [--fused]\n"); return 2; }
    CombinedSolverParameters params;
    for (int i = 2; i < argc; ++i) {
        if (!strcmp(argv[i], "--materialized")) params.useMaterializedJTJ = true;
        if (!strcmp(argv[i], "--fused")) params.useFusedJTJ = true;
    }
    FILE* f = fopen(argv[1], "rb");
    if (!f) { perror(argv[1]); return 2; }
    char magic[8];
    if (fread(magic, 1, 8, f) != 8 || memcmp(magic, "OPTB2REF", 8)) { fprintf(stderr, "bad magic\n"); return 2; }
    const std::string energy = rd_str(f), kind = rd_str(f);
    std::vector<unsigned int> dims(rd_u32(f));
    for (auto& d : dims) d = rd_u32(f);
    unsigned int nIterations = rd_u32(f), lIterations = rd_u32(f);
    const unsigned nparams = rd_u32(f);

    NamedParameters problemParams;
    std::vector<std::unique_ptr<float>> scalars;          // host scalars are passed by address and must outlive the solve
    std::vector<std::shared_ptr<OptImage>> unknowns;
    for (unsigned p = 0; p < nparams; ++p) {
        const unsigned type = rd_u32(f);
        const std::string name = rd_str(f);
        if (type == 1) {                                  // image: what createEmptyOptImage + update do in every CombinedSolver
            const bool isUnknown = rd_u32(f) != 0;
            const unsigned channels = rd_u32(f);
            std::vector<unsigned int> idims(rd_u32(f));
            size_t n = channels;
            for (auto& d : idims) { d = rd_u32(f); n *= d; }
            std::vector<float> host(n);
            if (fread(host.data(), 4, n, f) != n) return 2;
            auto im = createEmptyOptImage(idims, OptImage::Type::FLOAT, channels, OptImage::Location::GPU, isUnknown);
            im->update(host);
            problemParams.set(name, im);
            if (isUnknown) unknowns.push_back(im);
        } else if (type == 2) {
            scalars.emplace_back(new float);
            if (fread(scalars.back().get(), 4, 1, f) != 1) return 2;
            problemParams.set(name, (void*)scalars.back().get());
        } else if (type == 3) {                           // graph: OptGraph flattens to edgeCount, v0, v1, ... (NamedParameters.h:35-48)
            const unsigned arity = rd_u32(f), edges = rd_u32(f);
            std::vector<std::vector<int>> idx(arity, std::vector<int>(edges));
            for (auto& v : idx) if (fread(v.data(), 4, edges, f) != edges) return 2;
            problemParams.set(name, std::make_shared<OptGraph>(idx, OptImage::Location::GPU));
        } else { fprintf(stderr, "unknown parameter type %u\n", type); return 2; }
    }
    fclose(f);

    OptSolver solver(dims, energy, kind, params, /*doublePrecision=*/false, "backend_cuda", 1);
    NamedParameters solverParams;
    solverParams.set("nIterations", (void*)&nIterations);
    solverParams.set("lIterations", (void*)&lIterations);
    std::vector<SolverIteration> iters;
    const double cost = solver.solve(solverParams, problemParams, /*profiledSolve=*/true, iters);
    printf("harness: OptSolver::solve returned %.9g after %d recorded iterations\n", cost, (int)iters.size());
    // a checksum of the solved unknowns, so that the test can compare them with the Python-driven solve
    for (size_t u = 0; u < unknowns.size(); ++u) {
        std::vector<float> host(unknowns[u]->dataSize() / sizeof(float));
        unknowns[u]->copyTo(host);
        double s = 0, a = 0;
        for (float v : host) { s += v; a += v < 0 ? -v : v; }
        printf("harness: unknown %d sum=%.9g abs=%.9g\n", (int)u, s, a);
    }
    return 0;
}
