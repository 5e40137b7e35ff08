# FAST kernel: CTAs launched per resident CTA slot (1 = fully persistent ... 100000 = one cell per warp). Run on the GPU box.
for c in ${CELLS_LIST:-4}; do for m in ${MULT_LIST:-1 4 8 16 32 64 100000}; do
python bench.py --tune fast_cells=$c --tune fast_grid_mult=$m --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cells=$c mult=$m', round(d['value']), round(d['e2e']['value']), d['roofline']['stage_ms_per_step']['fast'])"
done; done
