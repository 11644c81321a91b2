# One bench line per BASELINE.json configuration (and solver-kind / precision variant) from ONE script, each with its
# roofline, cpu_baseline and parity fields: profiles/${TAG}_bench_<workload>[_variant].json.   TAG=r2n bash tools/bench_all.sh
set -x
cd ${GRAFT_REPO_ROOT:-.}
TAG=${TAG:-r2n}
OUT=gpurun_out
mkdir -p $OUT
run() {   # name, args...
  name=$1; shift
  timeout 900 python bench.py --steps 5 --warmup 3 --cpu-budget 10 "$@" 2> $OUT/${TAG}_$name.err | tail -1 > $OUT/${TAG}_bench_$name.json
  python - <<PY
import json
d = json.load(open("$OUT/${TAG}_bench_$name.json"))
r, p = d.get("roofline") or {}, d.get("parity") or {}
print("$name", "%.1f it/s" % d["value"], "e2e %.1f" % d["e2e"]["value"], "frac %s" % r.get("frac"), "nominal %s" % r.get("frac_of_nominal_8TBs"),
      "parity", p.get("ok"), p.get("rel_err_cost"), "cpu", (d.get("cpu_baseline") or {}).get("value"), d["clocks"]["reasons"])
PY
}
run volumetric_512
run poisson_4096 --workload poisson_4096
run poisson_4096_double --workload poisson_4096 --double
run poisson_4096_kindLMGPU --workload poisson_4096 --kind LMGPU
for w in optical_flow_2048 sfs_2048; do
  run $w --workload $w
  run ${w}_double --workload $w --double
  run ${w}_kindLMGPU --workload $w --kind LMGPU
  run ${w}_kindLMGPUdouble --workload $w --kind LMGPU --double
done
run arap_1m --workload arap_1m
run embedded_1m --workload embedded_1m
run cotangent_1m --workload cotangent_1m
run robust_1m --workload robust_1m
# the serial CPU row BASELINE.md section 4 promises (backend_cpu semantics: one thread), on the default workload's sample
run volumetric_512_cpu1 --cpu-threads 1 --steps 2 --warmup 1
run image_warping_512 --workload image_warping_512
run volumetric_256 --workload volumetric_256
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | tail -1 > $OUT/${TAG}_bench_reference_arm.json
cat $OUT/${TAG}_bench_reference_arm.json | cut -c1-300
# materialized-Jacobian modes: PCGStep1 through the CSR kernels, SpMV roofline in the line
run poisson_1024_materialized --workload poisson_4096 --size 1024 1024 --materialized 1 --no-cpu
run poisson_1024_fusedjtj --workload poisson_4096 --size 1024 1024 --materialized 2 --no-cpu
run arap_256k_fusedjtj --workload arap_1m --size 512 512 --materialized 2 --no-cpu
