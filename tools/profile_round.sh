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
# the first decode_megakernel launch of the command is the 8-step loop of bench.py's token check (right after the prefill)
ncu --set full --clock-control none --import-source on -k regex:decode_megakernel -s 0 -c 1 -o $out/${tag}_megakernel \
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
    -k regex:'gemm_tc|gemm_swapab|prefill_attn|norm_rows|rope_append|embed|argmax' python bench.py --steps 4 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
# Llama-3-8B: memory traffic of one persistent launch (for profiles/traffic.json), decode batches on the per-op path, prefill
ncu --set full --clock-control none -k regex:decode_megakernel -s 0 -c 1 -o $out/${tag}_megakernel_llama8b \
    python bench.py --workload llama3-8b --steps 8 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
ncu -i $out/${tag}_megakernel_llama8b.ncu-rep --page raw --csv > $out/${tag}_ncu_megakernel_llama8b_raw.csv 2>/dev/null
for b in 12 16 32 64; do
  python bench.py --workload llama3-8b --batch $b --steps 16 --no-cpu-baseline > $out/${tag}_bench_llama8b_b$b.json 2>/dev/null
done
python tools/prefill_probe.py llama3-8b 4096 4096 2>/dev/null | grep prefill > $out/${tag}_prefill_probe.txt
python tools/prefill_probe.py tinyllama-1.1b 512 512 2>/dev/null | grep prefill >> $out/${tag}_prefill_probe.txt
python tools/prefill_probe.py qwen2.5-14b 2048 2048 2>/dev/null | grep prefill >> $out/${tag}_prefill_probe.txt
python tools/gemm_bench.py > $out/${tag}_gemm_bench.txt 2>&1
rm -f $out/${tag}_megakernel_llama8b.ncu-rep
ls -la $out | tail -30
