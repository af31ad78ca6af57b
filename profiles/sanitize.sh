#!/bin/bash
# compute-sanitizer over one small-shape pass of every kernel family (tests/gpu_sanitize_pass.py), for the default kernels and
# for each alternative path-kernel variant; one GPU, run under gpurun.  Summaries land in gpurun_out/sanitize_*.log
run() {  # name tool env...   (compute-sanitizer answers "closed on this pool" here: profiles/r02_sanitizer_closed.log)
  name=$1; tool=$2; shift 2
  env "$@" timeout 1500 compute-sanitizer --tool $tool --print-limit 5 python tests/gpu_sanitize_pass.py > gpurun_out/sanitize_${name}_${tool}.log 2>&1
  echo "== $name $tool: exit $? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|sanitize pass' gpurun_out/sanitize_${name}_${tool}.log | tr '\n' ' ')"
}
run default memcheck X=1
run default racecheck X=1
run v3h2 memcheck CVR_PATH_KERNELS=v3 CVR_PATH_H_REV=2
run v3h2 racecheck CVR_PATH_KERNELS=v3 CVR_PATH_H_REV=2
run h3fwd memcheck CVR_PATH_H_FWD=3
run v2 memcheck CVR_PATH_KERNELS=v2
run generic memcheck CVR_PATH_KERNELS=generic
run cl1 memcheck CVR_PATHV_CL=1
run cl8 racecheck CVR_PATHV_CL=8
