show() { python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if 'yardstick' not in d['kernel']:
        print('%-42s %8.1f GB/s %s' % (d['kernel'], d['achieved_gbs'], d.get('frac_of_measured_peak', '')))
"; }
for u in 1 2; do echo "== DGS_FLAT_UNROLL=$u DGS_NORM_ELEMS=32768"; DGS_NORM_ELEMS=32768 DGS_FLAT_UNROLL=$u python tools/kernel_bench.py --mb 1024 --iters 20 | show | grep -v step; done
for u in 1 2 4; do echo "== DGS_STEP_UNROLL=$u"; DGS_STEP_UNROLL=$u python tools/kernel_bench.py --mb 1024 --iters 20 | show | grep step; done
echo "== small (64 MiB) arrays, FLAT_UNROLL 1 2 4 / STEP 1 2"
for u in 1 2 4; do DGS_FLAT_UNROLL=$u DGS_STEP_UNROLL=$u python tools/kernel_bench.py --mb 64 --iters 50 | show | sed "s/^/u=$u  /"; done
