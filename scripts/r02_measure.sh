#!/bin/bash
# End-of-round measurement set (round 2: same recipe as r01_measure.sh, more ncu columns) (one GPU): bench line, launch list of the same command, full captures of the trace kernel
# and of one path-tracer round per shade-kernel variant.  Outputs under gpurun_out/<tag>_* (reports are summarised on the
# box; only the small ones travel back -- gpurun_out/ is capped at 64 MiB).
tag=${1:-r02_s6}
what=${2:-all}
mkdir -p gpurun_out
if [ "$what" = all ]; then
python bench.py > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench_n1.err || exit 1
fi
python bench.py --steps 2 --warmup 3 --pt-iterations 4 --c5-spp 64 --office-spp 256 > gpurun_out/${tag}_bench_short.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --pt-iterations 4 --c5-spp 64 --office-spp 256 > gpurun_out/${tag}_ncu_launch.log 2>&1
python scripts/ncu_trace_target.py 2 > /dev/null || exit 1
ncu --set full --clock-control none --import-source on -k regex:trace_kernel -o gpurun_out/${tag}_trace_c3 -f \
    python scripts/ncu_trace_target.py 2 > gpurun_out/${tag}_ncu_trace.log 2>&1
python scripts/ncu_summary.py gpurun_out/${tag}_trace_c3.ncu-rep gpurun_out/${tag}_trace_c3_full_summary.csv > /dev/null
# one iteration of the diamond (lean shade kernel) and of C5 (extended shade kernel): warm-up off, first 40 launches
for fx in diamond many-point-lights; do
    FIXTURE=$fx WARM=0 python scripts/gpu_render_perf.py 1920 4 1 > /dev/null || exit 1
    FIXTURE=$fx WARM=0 ncu --set full --clock-control none -c 40 -o /tmp/${tag}_pt_$fx -f \
        python scripts/gpu_render_perf.py 1920 4 1 > gpurun_out/${tag}_ncu_pt_$fx.log 2>&1
    python scripts/ncu_summary.py /tmp/${tag}_pt_$fx.ncu-rep gpurun_out/${tag}_pt_${fx}_round_summary.csv > /dev/null
    FIXTURE=$fx WARM=0 ncu --set full --clock-control none --import-source on -k regex:shade_kernel -s 1 -c 1 -o gpurun_out/${tag}_shade_$fx -f \
        python scripts/gpu_render_perf.py 1920 4 1 >> gpurun_out/${tag}_ncu_pt_$fx.log 2>&1
done
du -sh gpurun_out; ls -la gpurun_out/
