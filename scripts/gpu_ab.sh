# GPU box: parity tests, then quick A/B of the pool variants at reduced spp (debug numbers)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | grep -v "^Loading" > gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log
q() { python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$1', round(d['value'],1), 'Msamples/s', round(d['mrays_per_s'],1), 'Mrays/s', {k:round(v,2) for k,v in (r.get('stage_ms_mean_over_ctas') or r.get('kernel_ms')).items()}, 'iters', r['wavefront_iterations'])"; }
for P in ${SLOTS:-512 768}; do
  FRT_WF_SLOTS=$P timeout 300 python bench.py --steps 2 --warmup 1 --spp ${SPP:-64} --no-e2e --no-cpu-baseline 2>gpurun_out/err_$P.log | tail -1 | q "wf P=$P" || tail -5 gpurun_out/err_$P.log
done
[ -n "$STREAMED" ] && timeout 300 python bench.py --steps 2 --warmup 1 --spp ${SPP:-64} --no-e2e --no-cpu-baseline --streamed-pool 2>/dev/null | tail -1 | q "streamed"
true
