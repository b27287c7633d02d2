#!/bin/bash
# Round 2, call U: compute-sanitizer over a subset of the parity tests (small sizes, every kernel family): memcheck with
# the affine rounds off (small inputs) and forced to 3, initcheck, then racecheck + synccheck on the shared-memory kernels.
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
SUB="tests/test_kat_gpu.py tests/test_zz_no_alloc.py"
KN='golden or chain_parallel or extended_domain_golden or kats or kc_msm_goldens or no_alloc or noalloc or (c_abi_prover_bit_exact and 18)'
run() { # name, tool, extra env, pytest args...
  name=$1; tool=$2; shift 2
  echo "== $name ($tool)"
  ( time timeout 420 $CS --tool $tool --error-exitcode 9 --print-limit 20 --log-file gpurun_out/r2_sanitizer_$name.log \
      python -m pytest -q -m gpu -x -p no:cacheprovider "$@" 2>&1 | tail -4 ) 2>&1 | grep -v "^$" | tail -8
  echo "rc=$? ; summary:"; grep -E "ERROR SUMMARY|Error|error" gpurun_out/r2_sanitizer_$name.log | sort | uniq -c | head -8
}
run memcheck_r0 memcheck $SUB tests/test_ntt_gpu.py tests/test_groth16_gpu.py -k "$KN"
B200_AFFINE_ROUNDS=3 run memcheck_r3 memcheck tests/test_kat_gpu.py tests/test_groth16_gpu.py -k "$KN"
B200_AFFINE_ROUNDS=2 run initcheck_r2 initcheck tests/test_kat_gpu.py tests/test_ntt_gpu.py tests/test_groth16_gpu.py -k "$KN"
run racecheck racecheck tests/test_kat_gpu.py tests/test_ntt_gpu.py -k "random_kats or extended_domain_golden or step_domain_golden"
run synccheck synccheck tests/test_kat_gpu.py tests/test_ntt_gpu.py -k "random_kats or extended_domain_golden or step_domain_golden"
ls -la gpurun_out/ | grep sanitizer
