# e2e of bench.py (DPVecEnv.step, pageable inputs) against the chunk size of the mapped host step
P='import sys,json; d=json.loads(sys.stdin.readlines()[-1]); e=d["e2e"]; print("e2e %.3e %.4f ms; pinned %.3e %.4f"%(e["value"],e["ms_per_step"],e["pinned"]["value"],e["pinned"]["ms_per_step"]))'
for c in ${CHUNKS:-512 768 1024 1536 2048 4096}; do echo -n "DM_HOST_CHUNK=$c ${EXTRA:-}: "; DM_HOST_CHUNK=$c python bench.py --no-cpu --no-sweep --steps 200 --warmup 20 ${EXTRA:-} 2>/dev/null | python -c "$P"; done
