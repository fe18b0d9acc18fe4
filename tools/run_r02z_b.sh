#!/bin/bash
# larger waves with a parking area to match (C2)
run() { env "$1" "$2" python bench.py --config c2 --steps 8 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('$1 $2 value %.4e ms %.3f primary %.3f' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_step']['primary']))"; }
run BF_WAVE_LANES=0 BF_PATH_STASH_MAX=65536
run BF_WAVE_LANES=67108864 BF_PATH_STASH_MAX=262144
run BF_WAVE_LANES=134217728 BF_PATH_STASH_MAX=524288
run BF_WAVE_LANES=16777216 BF_PATH_STASH_MAX=65536
