#!/bin/bash
# N-rank A/B of the round-2c switches on ONE box (weak 256^3 per GPU, short lines)
N=${1:-8}
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 10 --warmup 3 --no-strong1024 --no-e2e --no-cpu-baseline --no-parity --profile-steps 0 > gpurun_out/n${N}_$tag.json 2> gpurun_out/n${N}_$tag.err
  python - <<PY
import json
d = json.load(open("gpurun_out/n${N}_$tag.json"))
print("$tag", "%.3f ms" % d["ms_per_step"], d["pcg_iterations"][:3], d["config"].get("device_barriers_per_step"))
PY
}
shift
for v in "$@"; do if [ "$v" = default ]; then run default MF_X=1; else run "$v" "$v=1"; fi; done
