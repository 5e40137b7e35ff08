# FAST kernel layout: warp per cell with 1 / 2 / 4 / 8 cells per CTA. Run on the GPU box.
for c in 1 2 4 8; do
  python bench.py --tune fast_cells=$c --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cells=$c', round(d['value']), round(d['e2e']['value']), d['roofline']['stage_ms_per_step']['fast'])"
done
