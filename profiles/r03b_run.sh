# usage: bash profiles/r03b_run.sh -- 1.03 M-atom water box on one GPU (strong_scaling arm at N = 1) with and without the hot-row branch, clocks sampled
mkdir -p gpurun_out; TAG=r03b
ab() { n=$1; shift
  env "$@" timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_strong1_$n.json 2> gpurun_out/${TAG}_strong1_$n.err
  python -c "
import json
d=json.load(open('gpurun_out/${TAG}_strong1_$n.json'))['strong_scaling']; print('$n', round(d['ms_per_step'],2), d['clocks'], {k:(round(v['us_per_step']/1000,2)) for k,v in d['kernel_families'].items() if k.startswith('tn_msg')})
"
}
ab head TMDN_HOT_ROWS=0
ab hot1 TMDN_HOT_ROWS=1
ab head2 TMDN_HOT_ROWS=0
