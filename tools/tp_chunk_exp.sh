# TP decoder layer on all GPUs of the box: token-chunk shapes x all-reduce CTAs (dataflow schedule).
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
i=0
run() { i=$((i+1)); IREE_B200_COLLECTIVE_WATCHDOG_MS=30000 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+i)) bench.py --gpus $N --workload tp_decoder_layer --collective symmetric --reserved-sms $1 --ar-blocks $1 $2 --steps 150 --warmup 5 --e2e-steps 2 2>/dev/null | grep '^{' | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print(json.dumps({'tp_chunk_exp': {'world': d['n_gpus'], 'ar_blocks': $1, 'chunks': '$2', 'tflops': round(d['value'],1), 'ms_per_step': round(d['ms_per_step'],4)}}))" | tee -a gpurun_out/tp_chunk_exp_$N.jsonl; }
run 8 "--tp-chunks 4"
run 8 "--tp-chunks 2"
run 8 "--tp-chunks 3"
run 8 "--tp-chunk-rows 1024,3584,3584"
run 8 "--tp-chunk-rows 512,2560,5120"
run 8 "--tp-chunk-rows 1024,7168"
run 16 "--tp-chunk-rows 1024,3584,3584"
