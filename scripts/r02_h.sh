run() { tag=$1; shift; env "$@" python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/r02h_$tag.json 2>> gpurun_out/r02h.err; python - gpurun_out/r02h_$tag.json $tag <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    k=d["kernel_ms_per_frame"]
    print(sys.argv[2], round(d["ms_per_step"],2), "L0", round(k["k_hit"]+k["k_shadow"]+k["k_shade"],2), "L1", round(k["level1 k_hit"]+k["level1 k_shadow"]+k["level1 k_shade"],2), "deep", round(k["levels>=2 (hit+shadow+shade)"],2), "tail", round(k["k_tail"],2), d["image_sha256"][:8])
except Exception as e: print(sys.argv[2], "ERR", e)
PY
}
run base X=1
run tail1 B200RT_TAIL_LEVEL=1
run tail2 B200RT_TAIL_LEVEL=2
run tail4 B200RT_TAIL_LEVEL=4
run tail0 B200RT_TAIL_LEVEL=0
run deep4 B200RT_WAVES_DEEP=4
run deep16 B200RT_WAVES_DEEP=16
run deep32 B200RT_WAVES_DEEP=32
run w32 B200RT_WAVES=32
run w64 B200RT_WAVES=64
tail -3 gpurun_out/r02h.err
