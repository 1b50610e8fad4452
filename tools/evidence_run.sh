#!/bin/bash
# Evidence run on N GPUs of one box (round-2 numbers quoted in BASELINE.md / profiles/README.md):
#   tools/evidence_run.sh N [tests]     under   gpurun --gpus N -- 'bash tools/evidence_run.sh N'
# Everything lands in gpurun_out/ev_nN/.  Pass "tests" as second argument to also run the multi-GPU pytest file.
N=${1:-1}
OUT=gpurun_out/ev_n$N
mkdir -p $OUT
if [ "$N" -gt 1 ]; then
  T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
else
  T="timeout 600 python"
fi
port=29600
run() { port=$((port+1)); if [ "$N" -gt 1 ]; then $T --master-port $port "$@"; else $T "$@"; fi; }
jl() { grep "^{" | tail -1; }
# headline + its in-process library baseline + fp32-accurate arm
run bench.py --gpus $N --steps 200 --warmup 20 2>$OUT/bench.err | jl > $OUT/bench_cifar_eamsgd.json
# config 2: MNIST-shaped CNN, Downpour vs EASGD; the synchronous all-reduce trainer
run bench.py --gpus $N --steps 200 --warmup 20 --model mnist --optim downpour --no-fp32-arm 2>>$OUT/bench.err | jl > $OUT/bench_mnist_downpour.json
run bench.py --gpus $N --steps 200 --warmup 20 --model mnist --optim easgd --no-fp32-arm 2>>$OUT/bench.err | jl > $OUT/bench_mnist_easgd.json
run bench.py --gpus $N --steps 200 --warmup 20 --optim sync_eamsgd --no-baseline --no-fp32-arm 2>>$OUT/bench.err | jl > $OUT/bench_cifar_sync_eamsgd.json
# config 4: BiCNN with server-side Adam on the asynchronous device parameter server vs the NCCL formulation
run tools/bench_bicnn.py --gpus $N --steps 100 --warmup 10 2>>$OUT/bench.err | jl > $OUT/bicnn_mpit.json
run tools/bench_bicnn.py --gpus $N --steps 30 --warmup 5 --impl torch 2>>$OUT/bench.err | jl > $OUT/bicnn_torch.json
if [ "$N" -gt 1 ]; then
  # config 5: exchange sweep 1 KB .. 1 GiB with the NCCL formulation next to it
  run -m mpit_b200.launch.ptest --nccl --iters 5 --out $OUT/sweep.jsonl > $OUT/sweep.log 2>&1
  run tools/bench_pulled_linear.py $OUT/pulled_linear.jsonl > $OUT/pulled_linear.log 2>&1
  run -m mpit_b200.launch.ptest2 --device cuda --rounds 8 --ssize 3458600 > $OUT/ptest2_device.log 2>&1
fi
if [ "$2" = "tests" ]; then
  if [ "$N" -gt 1 ]; then python -m pytest tests/test_gpu_multi.py -q -m gpu > $OUT/pytest_multi.log 2>&1
  else python -m pytest tests -q -m gpu > $OUT/pytest_gpu.log 2>&1; fi
  tail -3 $OUT/pytest_*.log
fi
for f in $OUT/*.json; do echo "== $f"; python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read())
    keys = ("value", "ms_per_step", "vs_baseline", "n_gpus")
    print({k: d.get(k) for k in keys}, {k: (d.get("e2e") or {}).get(k) for k in ("value",)}, (d.get("fp32_accurate") or {}).get("value"))
except Exception as e:
    print("unreadable:", e)
PY
done
ls $OUT
