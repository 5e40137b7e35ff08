# e2e pipeline sweep: compute streams x chunks x first-chunk split (run on the GPU box)
for cfg in "2 8 2" "3 8 2" "4 8 2" "3 12 2" "4 12 2" "3 16 2" "4 16 2" "4 16 4" "3 8 4" "4 24 2" "4 32 2"; do set -- $cfg
  python bench.py --tune e2e_streams=$1 --tune e2e_chunks=$2 --tune e2e_first_div=$3 --steps 8 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('streams $1 chunks $2 firstdiv $3 e2e', round(d['e2e']['value']), 'value', round(d['value']))"
done
