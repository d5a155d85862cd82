#!/bin/bash
# Multi-GPU evidence of one 8-GPU box, builder-run (gpurun --gpus 8 -- bash tools/scale_run.sh):
#   gpurun_out/r02_scale_4k.jsonl    bench.py --config 4k at N = 1, 2, 4, 8 (one process per GPU, torchrun)
#   gpurun_out/r02_pcie.jsonl        tools/pcie_probe.py at N = 1, 2, 4, 8
#   gpurun_out/r02_mgpu_e2e.json     tools/mgpu_e2e.py (single process, of_mgpu_*)
#   gpurun_out/r02_mgpu_pytest.log   the GPU tests that need >= 2 GPUs
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
: > gpurun_out/r02_scale_4k.jsonl; : > gpurun_out/r02_pcie.jsonl
python -m pytest tests/test_gpu_round2.py -m gpu -q -k "mgpu or two_devices or torchrun or cpp_caller" 2>&1 | tail -5 > gpurun_out/r02_mgpu_pytest.log
port=29700
for n in 1 2 4 8; do
  port=$((port+1))
  if [ $n = 1 ]; then
    python bench.py --config 4k --gpus 1 --steps 10 --warmup 3 --lean >> gpurun_out/r02_scale_4k.jsonl 2>> gpurun_out/r02_scale.err
    python tools/pcie_probe.py >> gpurun_out/r02_pcie.jsonl 2>> gpurun_out/r02_scale.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --config 4k --gpus $n --steps 10 --warmup 3 --lean >> gpurun_out/r02_scale_4k.jsonl 2>> gpurun_out/r02_scale.err
    port=$((port+1))
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port tools/pcie_probe.py >> gpurun_out/r02_pcie.jsonl 2>> gpurun_out/r02_scale.err
  fi
done
python tools/mgpu_e2e.py > gpurun_out/r02_mgpu_e2e.json 2>> gpurun_out/r02_scale.err
tail -5 gpurun_out/r02_scale.err
