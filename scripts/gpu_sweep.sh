mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -3
for leaf in 4 8 16; do
  FRT_MAX_LEAF=$leaf python bench.py --steps 1 --warmup 1 --spp 128 --no-e2e --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('leaf $leaf', round(d['value'],1), d['roofline']['kernel_ms'], round(d['roofline']['nodes_per_ray'],2), round(d['roofline']['tris_per_ray'],2))"
done
