# FAST change check: parity subset + device-resident bench with the stage times
python -m pytest tests/test_gpu_parity.py tests/test_reference_pin.py -m gpu -x -q 2>&1 | tail -1
for c in ${CELLS_LIST:-4}; do
python bench.py --tune fast_cells=$c --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cells=$c', round(d['value']), round(d['e2e']['value']), d['roofline']['stage_ms_per_step'])"
done
