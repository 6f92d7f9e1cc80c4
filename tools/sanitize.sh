#!/bin/bash
# compute-sanitizer pass over the library's kernels: memcheck, racecheck, initcheck and synccheck on
# smoke() and on the small-recipe GPU parity tests.  Only kernels of libpigs_b200.so are checked (namespace pigs ->
# "4pigs" in the mangled name); torch's own kernels are skipped.
# usage: [TOOLS="memcheck racecheck"] tools/sanitize.sh <tag> [seconds per invocation]  -> gpurun_out/<tag>/sanitizer_*.log
set -u
TAG=${1:-rXX}
CAP=${2:-420}
OUT=gpurun_out/$TAG
mkdir -p $OUT
SEL='synth2d or torus or wide or agg_model or culled'
TESTS="tests/test_gpu_parity.py tests/test_gpu_aggregate.py tests/test_culled.py"
for tool in ${TOOLS:-memcheck racecheck initcheck synccheck}; do
  extra=""
  [ $tool = initcheck ] && extra="--check-api-memory-access no"
  [ $tool = racecheck ] && extra="--racecheck-report all"
  timeout $CAP compute-sanitizer --tool $tool $extra --kernel-name kns=4pigs --error-exitcode 7 --print-limit 40 \
      --log-file $OUT/sanitizer_${tool}_smoke.log python -c "import __graft_entry__ as g; g.smoke()" \
      > $OUT/sanitizer_${tool}_smoke.out 2>&1
  echo "$tool smoke exit $?" | tee -a $OUT/sanitizer_summary.txt
  timeout $CAP compute-sanitizer --tool $tool $extra --kernel-name kns=4pigs --error-exitcode 7 --print-limit 40 \
      --log-file $OUT/sanitizer_${tool}_tests.log python -m pytest $TESTS -m gpu -q -x -k "$SEL" -p no:cacheprovider \
      > $OUT/sanitizer_${tool}_tests.out 2>&1
  echo "$tool tests exit $? ($(tail -1 $OUT/sanitizer_${tool}_tests.out))" | tee -a $OUT/sanitizer_summary.txt
  grep -h "ERROR SUMMARY\|RACECHECK SUMMARY" $OUT/sanitizer_${tool}_smoke.log $OUT/sanitizer_${tool}_tests.log | tee -a $OUT/sanitizer_summary.txt
done
