#!/bin/bash
# Dev tool: C1 step (1-D, 1K x 10K) against the work-item granularity of the small-problem mode (pack.cu small_min_stream).
mkdir -p gpurun_out
out=gpurun_out/${1:-r04c}_small_sweep.txt
: > $out
run() {
    python bench.py --workload C1 --steps 50 --warmup 10 --no-legs --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$1', 'ms/step %.4f' % d['ms_per_step'], 'launches/step %.1f' % (d['gpu_launches']/d['steps']), 'fwd %.4f bwd %.4f other %.4f' % (d['kernel_ms']['eval_fwd'], d['kernel_ms']['eval_bwd'], d['kernel_ms']['other_per_step']), 'e2e %.4f' % d['e2e']['ms_per_step'])
" >> $out
}
PIGS_SMALL_PAIRS=0 run "three-class"
for f in 8 16 32 64 128; do
    for b in 16 32 64 128; do
        PIGS_SMALL_MIN_STREAM_FWD=$f PIGS_SMALL_MIN_STREAM_BWD=$b run "single fwd_min=$f bwd_min=$b"
    done
done
cat $out
