# G1 launch shapes after the packed-arithmetic changes
set -u
mkdir -p gpurun_out
one() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%-52s kernel %8.2f us frac %.3f' % (sys.argv[1], d['roofline']['kernel_ms']*1e3, d['roofline']['frac']))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
}
R1="python bench.py --gpus 1 --no-cpu --no-e2e --no-sweep --no-configs --warmup 3 --steps 20"
for e in 131072 262144 524288 1048576; do
for rb in "g1c:--robot unitree_g1 --combined" "g1:--robot unitree_g1"; do
name=${rb%%:*}; args=${rb#*:}
timeout 300 $R1 --envs $e $args > gpurun_out/g.json 2>gpurun_out/g.err; one "$name $e default" gpurun_out/g.json
DM_WIDE=2 timeout 300 $R1 --envs $e $args > gpurun_out/g.json 2>gpurun_out/g.err; one "$name $e WIDE=2" gpurun_out/g.json
DM_WIDE=0 timeout 300 $R1 --envs $e $args > gpurun_out/g.json 2>gpurun_out/g.err; one "$name $e WIDE=0" gpurun_out/g.json
done
done
for e in 131072 262144; do
timeout 300 $R1 --envs $e --layout soa --robot unitree_g1 --combined > gpurun_out/g.json 2>gpurun_out/g.err; one "g1c soa $e" gpurun_out/g.json
timeout 300 $R1 --envs $e --layout soa > gpurun_out/g.json 2>gpurun_out/g.err; one "h soa $e" gpurun_out/g.json
done
