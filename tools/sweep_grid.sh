for m in 2 4 8 16 32 64 256; do echo "== DGS_GRID_MULT=$m"; DGS_GRID_MULT=$m python tools/kernel_bench.py --mb 1024 --iters 20 | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print('%-42s %8.1f GB/s %s' % (d['kernel'], d['achieved_gbs'], d.get('frac_of_measured_peak', '')))
"; done
