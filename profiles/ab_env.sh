# A/B of one environment knob on a product arm, interleaved on one box:
#   bash profiles/ab_env.sh VAR val_a val_b [steps] [resnet50|wide_deep]
V=$1; A=$2; B=$3; S=${4:-40}; W=${5:-resnet50}
for x in $A $B $A $B; do
  env $V=$x python bench.py --workload $W --arms primary --steps $S --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$W $V=$x', d['ms_per_step'], 'ms', d['value'], d['unit'], d['clocks']['sm_mhz'])"
done
