// TEST INFRASTRUCTURE ONLY -- C ABI of the CPU oracle (see oracle/README.md, opt_oracle.h).
#include "opt_oracle.h"
#include "energies.hpp"
#include <string>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace oracle;

struct OptOracle_State { Opt_InitializationParameters ip; };
struct OptOracle_Problem { std::string name; bool lm; };
struct OptOracle_Plan { PlanBase* impl; };

static std::string basename_noext(const char* filename) {
    std::string s(filename);
    size_t slash = s.find_last_of("/\\");
    if (slash != std::string::npos) s = s.substr(slash + 1);
    size_t dot = s.rfind(".t");
    if (dot != std::string::npos && dot + 2 == s.size()) s = s.substr(0, dot);
    return s;
}

template <template <class> class E>
static PlanBase* make(bool dbl, const unsigned* dims, bool lm) {
    if (dbl) return new Solver<E<double>>(dims, lm);
    return new Solver<E<float>>(dims, lm);
}

static PlanBase* create(const std::string& n, bool dbl, const unsigned* dims, bool lm) {
#define ORACLE_CASE(E) if (n == E<float>::name()) return make<E>(dbl, dims, lm);
    ORACLE_ENERGY_LIST(ORACLE_CASE)
#undef ORACLE_CASE
    return nullptr;
}

extern "C" {

OptOracle_State* OptOracle_NewState(Opt_InitializationParameters params) { return new OptOracle_State{params}; }
void OptOracle_FreeState(OptOracle_State* s) { delete s; }

OptOracle_Problem* OptOracle_ProblemDefine(OptOracle_State*, const char* filename, const char* solverkind) {
    std::string n = basename_noext(filename);
    unsigned one[3] = {1, 1, 1};
    PlanBase* probe = create(n, false, one, false);
    if (!probe) { fprintf(stderr, "OptOracle: energy '%s' is not one of the shipped examples\n", filename); return nullptr; }
    delete probe;
    return new OptOracle_Problem{n, std::string(solverkind).find("LM") != std::string::npos};   // o.t:474
}
void OptOracle_ProblemDelete(OptOracle_State*, OptOracle_Problem* p) { delete p; }

OptOracle_Plan* OptOracle_ProblemPlan(OptOracle_State* s, OptOracle_Problem* p, unsigned int* dims) {
    PlanBase* impl = create(p->name, s->ip.doublePrecision != 0, dims, p->lm);
    impl->verbosity = s->ip.verbosityLevel;
    impl->threads = s->ip.numthreads > 0 ? s->ip.numthreads : 1;
    return new OptOracle_Plan{impl};
}
void OptOracle_PlanFree(OptOracle_State*, OptOracle_Plan* pl) { delete pl->impl; delete pl; }

void OptOracle_SetSolverParameter(OptOracle_State* s, OptOracle_Plan* pl, const char* name, void* value) {
    if (!pl->impl->sp.set(name, value) && s->ip.verbosityLevel > 0)
        printf("Warning: tried to set nonexistent solver parameter %s\n", name);
}
void OptOracle_ProblemInit(OptOracle_State*, OptOracle_Plan* pl, void** pp) { pl->impl->init(pp); }
int OptOracle_ProblemStep(OptOracle_State*, OptOracle_Plan* pl, void** pp) { return pl->impl->step(pp); }
void OptOracle_ProblemSolve(OptOracle_State* s, OptOracle_Plan* pl, void** pp) {       // o.t:3328-3333
    OptOracle_ProblemInit(s, pl, pp);
    while (OptOracle_ProblemStep(s, pl, pp)) {}
}
double OptOracle_ProblemCurrentCost(OptOracle_State*, OptOracle_Plan* pl) { return pl->impl->cost(); }

long OptOracle_UnknownCount(OptOracle_Plan* pl) { return pl->impl->unknown_count(); }
long OptOracle_PcgIterations(OptOracle_Plan* pl) { return pl->impl->pcg_iterations; }
int OptOracle_CostHistory(OptOracle_Plan* pl, double* out, int cap) {
    int n = (int)pl->impl->cost_history.size();
    for (int i = 0; i < n && i < cap; ++i) out[i] = pl->impl->cost_history[i];
    return n;
}
void OptOracle_ApplyJTJ(OptOracle_Plan* pl, void** pp, const void* p, void* Ap) { pl->impl->debug_apply_jtj(pp, p, Ap); }
void OptOracle_EvalJTF(OptOracle_Plan* pl, void** pp, void* g, void* pre) { pl->impl->debug_eval_jtf(pp, g, pre); }

}  // extern "C"
