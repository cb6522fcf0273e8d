#!/bin/bash
# Profile recipe of a round (run on the GPU box through gpurun): bench line first (no profiler), then the ncu launch
# list of the same command, then one full capture of the persistent decode kernel, then the per-phase trace.
# Usage: tools/profile_round.sh <tag>     -> gpurun_out/<tag>_*
set -u
tag=${1:-rX}
out=gpurun_out
mkdir -p $out
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_8step.json 2> $out/${tag}_bench_8step.err || exit 1
python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $out/${tag}_ncu_launches.csv \
    python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_launches.log 2>&1
# the timed launch (8 steps) is the second decode_megakernel launch of that command with more than one step:
# launches in order: prefill-side single-step launches (none for a 512-row chunk), warm-up loop (3 steps), timed loop (8)
ncu --set full --clock-control none --import-source on -k regex:decode_megakernel -s 1 -c 1 -o $out/${tag}_megakernel \
    python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_full.log 2>&1
ncu -i $out/${tag}_megakernel.ncu-rep --page raw --csv > $out/${tag}_ncu_megakernel_raw.csv 2>/dev/null
DLLM_MK_TRACE=$out/${tag}_phase_trace.csv python bench.py --steps 6 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
python tools/mk_trace_summary.py $out/${tag}_phase_trace.csv > $out/${tag}_phase_trace_summary.txt
# Llama-3-8B on the same kernel, batch 1 and 8
for b in 1 8; do
  python bench.py --workload llama3-8b --batch $b --steps 32 --no-cpu-baseline > $out/${tag}_bench_llama8b_b$b.json 2>/dev/null
done
DLLM_MK_TRACE=$out/${tag}_phase_trace_llama8b.csv python bench.py --workload llama3-8b --steps 6 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
python tools/mk_trace_summary.py $out/${tag}_phase_trace_llama8b.csv > $out/${tag}_phase_trace_llama8b_summary.txt
# prefill launch list (tcgen05 GEMMs + flash attention): one 512-token chunk
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_ncu_prefill_launches.csv \
    -k regex:'gemm_tc|prefill_attn|norm_rows|rope_append|embed' python bench.py --steps 4 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
ls -la $out | tail -20
