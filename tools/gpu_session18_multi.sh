#!/bin/bash
# multi-GPU session (gpurun --gpus N): the single-process group entries, then the strong-scaling bench at N ranks (and N/2 when N >= 4)
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > $O/s18_smi_$N.txt
timeout 600 python -m pytest tests/test_gpu_group.py -m gpu -q -p no:cacheprovider > $O/s18_group_$N.log 2>&1; echo "group tests exit $?"; tail -3 $O/s18_group_$N.log
bench() { local n=$1; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n > $O/r2_bench_c2_${n}gpu_v5.json 2> $O/r2_bench_c2_${n}gpu_v5.err; echo "bench $n gpus exit $?"; }
bench $N
if [ $N -ge 4 ]; then bench $((N / 2)); fi
python - $N <<'P'
import json, sys
n = int(sys.argv[1])
for k in sorted({n, n // 2 if n >= 4 else n}):
    f = f"gpurun_out/r2_bench_c2_{k}gpu_v5"
    try:
        d = json.load(open(f + ".json"))
        print(k, "GPUs", d["scaling"], round(d["value"]), "signals/s", round(d["ms_per_step"], 2), "ms/step; e2e", round(d["e2e"]["value"]) if d.get("e2e") else None, "; weak", (d.get("weak") or {}).get("value"), "; gathered", d.get("gathered"), "; step_ms", d.get("step_ms"))
    except Exception as e:
        print(k, "FAILED", e, open(f + ".err").read()[-1500:])
P
