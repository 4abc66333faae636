#!/bin/bash
# round 2, two-GPU session: the single-process group entry (ncclCommInitAll) and the strong-scaling bench over 2 ranks
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/s5_smi.txt
timeout 600 python -m pytest tests/test_gpu_group.py tests/test_gpu_irls.py -m gpu -q -s -p no:cacheprovider -k "group or c4" > gpurun_out/s5_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/s5_pytest.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/s5_bench_c2_2gpu.json 2> gpurun_out/s5_bench_c2_2gpu.err; echo "bench 2gpu exit $?"
python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/s5_bench_c2_1gpu.json 2> gpurun_out/s5_bench_c2_1gpu.err; echo "bench 1gpu exit $?"
python - <<'P'
import json
for f in ("s5_bench_c2_1gpu", "s5_bench_c2_2gpu"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, d["n_gpus"], d["scaling"], round(d["value"]), "signals/s", round(d["ms_per_step"], 2), "ms/step e2e", d["e2e"]["value"] if d["e2e"] else None, "weak", d.get("weak"), "gathered", d.get("gathered"))
    except Exception as e:
        print(f, "FAILED", e, open(f"gpurun_out/{f}.err").read()[-1500:])
P
