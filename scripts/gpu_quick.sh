#!/bin/bash
# quick single-GPU bench lines: gpu_quick.sh <outdir> "scale dtype [ENV=..]" ...
export PYTHONUNBUFFERED=1
O=gpurun_out/$1; shift; mkdir -p $O
for cfg in "$@"; do set -- $cfg; sc=$1; dt=$2; shift 2
  tag=$(echo "$*" | tr ' =/' '___'); [ -z "$tag" ] && tag=def
  env "$@" B200_BAND_VERBOSE=1 timeout 600 python bench.py --scale $sc --dtype $dt --steps 20 --warmup 3 --no-e2e --no-cpu --no-parity > $O/b_s${sc}_${dt}_$tag.json 2> $O/b_s${sc}_${dt}_$tag.err
  python -c "
import json
d=json.loads([l for l in open('$O/b_s${sc}_${dt}_$tag.json') if l.startswith('{')][-1]); print('s$sc $dt $tag: GTEPS', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],3))" || tail -3 $O/b_s${sc}_${dt}_$tag.err
done
echo "elapsed $SECONDS"
