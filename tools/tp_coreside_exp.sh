# TP layer experiments on whatever GPUs the box has: schedule x SM reservation x all-reduce shape.
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
i=0
run() { i=$((i+1)); IREE_B200_COLLECTIVE_WATCHDOG_MS=30000 IREE_B200_ALL_REDUCE_TUNING="$1" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+i)) bench.py --gpus $N --workload tp_decoder_layer --collective ${6:-symmetric} --tp-schedule $5 --reserved-sms $2 --ar-blocks $3 --tp-chunks $4 --steps 100 --warmup 5 --e2e-steps 2 2>/dev/null | grep '^{' | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print(json.dumps({'tp_exp': {'world': d['n_gpus'], 'collective': '${6:-symmetric}', 'tuning': '$1', 'reserved_sms': $2, 'ar_blocks': $3, 'chunks': $4, 'schedule': '$5', 'tflops': round(d['value'],1), 'ms_per_step': round(d['ms_per_step'],4)}}))" | tee -a gpurun_out/tp_exp_$N.jsonl; }
run 1024,4,0 16 16 4 dataflow
run 1024,4,0 16 16 4 staged
run 1024,4,0 16 16 4 dataflow nccl
run 512,4,0 0 64 4 dataflow
run 512,2,1 8 64 4 dataflow
run 1024,2,1 32 32 4 dataflow
run 1024,4,0 16 16 8 dataflow
run 1024,2,1 24 24 4 dataflow
