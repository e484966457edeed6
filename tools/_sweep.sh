python -m pytest tests -m gpu -x -q 2>&1 | tail -2
run() { timeout 300 python bench.py --steps 3 --warmup 1 --no-cpu-baseline "$@" 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), end=' ')"; }
echo -n "book1: "; run --spp 200; echo
for sc in 0 6 7; do echo -n "scene$sc: "; run --scene $sc --width 600 --spp 100; echo; done
