# GPU box, round 2 call H: ncu evidence.  Launch list of the bench command, the DRAM bytes of the full-size launch, and
# full captures of the three dominant kernels (Cornell path kernel, teapot path kernel, soup ray cast).  The .ncu-rep
# files stay in /tmp on the box (three of them exceed gpurun_out's 64 MiB): they are summarised here into text.
set -x; mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --spp 8 --no-e2e --no-cpu-baseline --no-roofline --no-configs"
$B > gpurun_out/prof_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/prof_ncu1.log 2>&1
bash scripts/gpu_traffic.sh
prof() { # name, kernel regex, command...
  name=$1; kern=$2; shift 2
  "$@" > gpurun_out/prof_${name}_plain.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$kern --launch-skip 1 --launch-count 1 -f -o /tmp/prof_$name "$@" > gpurun_out/prof_${name}_ncu.log 2>&1
  { echo "# ncu --set full --clock-control none --import-source on -k regex:$kern --launch-skip 1 --launch-count 1 $*"
    python scripts/ncu_summary.py /tmp/prof_$name.ncu-rep; echo; echo "# per source line (scripts/ncu_lines.py)"
    python scripts/ncu_lines.py /tmp/prof_$name.ncu-rep 40; } > gpurun_out/ncu_$name.txt 2>&1
}
prof wavefront k_wavefront python scripts/prof_target.py cornell 8
prof teapot k_wavefront python scripts/prof_target.py teapot 8
prof soup_trace k_trace python scripts/prof_target.py soup 1e7 4194304
PROF_ANY=1 prof soup_trace_any k_trace python scripts/prof_target.py soup 1e7 4194304
NCU_METRICS_OUT=gpurun_out/ncu_metrics.json python scripts/ncu_metrics.py \
  "k_wavefront=/tmp/prof_wavefront.ncu-rep:1080p x 8 spp Cornell (scripts/gpu_round2_h.sh)" \
  "k_wavefront.traffic=gpurun_out/traffic_1024.csv:1080p x 1024 spp Cornell, the launch bench.py times" \
  "k_wavefront_teapot=/tmp/prof_teapot.ncu-rep:1080p x 8 spp teapot" \
  "k_trace=/tmp/prof_soup_trace.ncu-rep:10^7-triangle soup, 2^22 closest-hit rays" \
  "k_trace_any=/tmp/prof_soup_trace_any.ncu-rep:10^7-triangle soup, 2^22 occlusion rays" > /dev/null
ls -la gpurun_out | tail -20
