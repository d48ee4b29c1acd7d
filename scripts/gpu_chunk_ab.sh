set -u
OUT=gpurun_out; mkdir -p $OUT
run() { tag=$1; shift; env "$@" timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --pipeline-chunks $CH > $OUT/r01l_$tag.json 2>$OUT/r01l_$tag.err; python - $OUT/r01l_$tag.json $tag <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[2], "value %.0f q/s  ms/step %.3f  e2e %.0f  clocks %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"]["sm_mhz"]))
PY
}
CH=4 run c4 EVI_MIN_CHUNK=16384
CH=16 run c16 EVI_MIN_CHUNK=16384
CH=40 run c40_16k EVI_MIN_CHUNK=16384
CH=80 run c80_8k EVI_MIN_CHUNK=8192
CH=110 run c110_6k EVI_MIN_CHUNK=6144
CH=160 run c160_4k EVI_MIN_CHUNK=4096
CH=2 run c2 EVI_MIN_CHUNK=16384
CH=4 run c4_again EVI_MIN_CHUNK=16384
