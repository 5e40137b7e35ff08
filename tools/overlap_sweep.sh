for cfg in "0 2 2" "5 2 2" "4 2 2" "5 4 2" "5 4 3" "4 4 3" "5 6 3" "0 4 3" "4 8 4"; do set -- $cfg
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-latency --tune fast_ctas=$1 --tune run_chunks=$2 --tune run_streams=$3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('fast_ctas=$1 chunks=$2 streams=$3', round(d['value']), d['roofline']['stage_ms_per_step']['fast'])"
done
