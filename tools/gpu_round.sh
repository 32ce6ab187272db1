#!/bin/bash
# One GPU-box visit: parity tests, smoke, the bench lines and the ncu evidence, each stage
# under its own timeout and logged to gpurun_out/<tag>_*.  Usage: tools/gpu_round.sh TAG [stages...]
TAG=${1:-r1}; shift
STAGES=${@:-tests smoke bench ncu}
O=gpurun_out; mkdir -p $O
run() { name=$1; shift; local t=$1; shift; echo "== $name"; timeout $t "$@" > $O/${TAG}_$name.log 2>&1; echo "   exit $? ($(tail -c 300 $O/${TAG}_$name.log | tr '\n' ' ' | tail -c 200))"; }
for st in $STAGES; do case $st in
tests) run pytest 900 python -m pytest tests -m gpu -x -q ;;
newtests) run pytest_new 600 python -m pytest tests -m gpu -x -q -k "pipelined or decode or selection or full_size" ;;
pending)  # code written without GPU time: first thing to run next round
  IREE_B200_PENDING_VALIDATION=1 run pytest_pending 300 python -m pytest tests/test_zz_mxfp4_gpu.py tests/test_zz_pipelined_bias_gpu.py tests/test_zz_mmt4d_gpu.py -m gpu -q
  run pending_bench 300 python tools/pending_bench.py --out $O/${TAG}_pending_bench.json
  run bench_linear_bf16 300 python bench.py --workload linear_bf16 --no-cpu-baseline --steps 100 ;;
smoke) run smoke 300 python -c "import __graft_entry__ as g; g.smoke()" ;;
boundary)  # the drop-in boundary through the reference's own runtime and stock tools
  run pytest_boundary 400 python -m pytest tests/test_ref_runtime.py tests/test_vmfb.py tests/test_cabi.py -m gpu -x -q
  run ref_runtime_bench 600 python tools/ref_runtime_bench.py
  run vmfb_bench 700 python tools/vmfb_bench.py
  run mmt4d_bench 300 python tools/mmt4d_bench.py --min-ms 150 ;;
bench)
  run bench_matmul_bf16 400 python bench.py
  for w in matmul_i8 batch_matmul attention_prefill attention_prefill_bf16 attention_decode; do
    run bench_$w 300 python bench.py --workload $w --no-cpu-baseline --steps 100
  done
  run bench_reference 300 python bench.py --impl reference --steps 2 --warmup 1 ;;
final)
  run pytest 900 python -m pytest tests -m gpu -q
  run smoke 300 python -c "import __graft_entry__ as g; g.smoke()"
  run bench_matmul_bf16 400 python bench.py
  run bench_attention_prefill 300 python bench.py --workload attention_prefill --no-cpu-baseline --steps 200
  run bench_attention_prefill_bf16 300 python bench.py --workload attention_prefill_bf16 --no-cpu-baseline --steps 200
  run bench_reference 300 python bench.py --impl reference --steps 2 --warmup 1 ;;
prefillsplit)
  run pytest_attn 400 python -m pytest tests/test_attention_gpu.py -m gpu -x -q
  run bench_prefill_split 200 python bench.py --workload attention_prefill --no-cpu-baseline --steps 200 --e2e-steps 1
  run bench_prefill_nosplit 200 python bench.py --workload attention_prefill --no-cpu-baseline --steps 200 --e2e-steps 1 --attention-flags 0x10000
  run bench_prefill_bf16_split 200 python bench.py --workload attention_prefill_bf16 --no-cpu-baseline --steps 200 --e2e-steps 1 ;;
prefillsweep)
  run pytest_prefill 300 python -m pytest tests/test_attention_gpu.py -m gpu -x -q -k "prefill"
  for v in 0x2000 0x4000 0x8000; do
    run bench_prefill_$v 200 python bench.py --workload attention_prefill --no-cpu-baseline --steps 200 --attention-flags $v --e2e-steps 1
  done
  run bench_prefill_bf16_0x8000 200 python bench.py --workload attention_prefill_bf16 --no-cpu-baseline --steps 200 --attention-flags 0x8000 --e2e-steps 1 ;;
ncufinal)
  run ncu_prefill 300 ncu --set full --clock-control none --import-source on -k regex:attention_prefill -c 1 -o $O/${TAG}_attn_prefill -f python bench.py --workload attention_prefill --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1
  run ncu_f8 300 ncu --set full --clock-control none --import-source on -k regex:matmul_f8e4m3fn_f32 -c 1 -o $O/${TAG}_gemm_f8 -f python bench.py --workload matmul_f8 --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1
  run ncu_i8 300 ncu --set full --clock-control none --import-source on -k regex:matmul_i8_i32 -c 1 -o $O/${TAG}_gemm_i8 -f python bench.py --workload matmul_i8 --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 ;;
f8) run bench_matmul_f8 300 python bench.py --workload matmul_f8 --no-cpu-baseline --steps 100 ;;
decode) run bench_attention_decode 300 python bench.py --workload attention_decode --no-cpu-baseline --steps 200 ;;
ncu)
  run ncu_launches 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_matmul_bf16.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline
  run ncu_decode 400 ncu --set full --clock-control none --import-source on -k regex:decode_cluster -c 1 -o $O/${TAG}_attn_decode_cluster -f python bench.py --workload attention_decode --steps 3 --warmup 3 --no-cpu-baseline ;;
symcheck)
  N=$(nvidia-smi -L | wc -l)
  run symcheck 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 tools/symmetric_check.py ;;
symcheck64)
  N=$(nvidia-smi -L | wc -l)
  run symcheck64 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 tools/symmetric_check.py --max-blocks 64 ;;
tp)
  N=$(nvidia-smi -L | wc -l)
  for c in symmetric nccl; do
    run bench_tp_${c}_n$N 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $N --workload tp_decoder_layer --collective $c --steps 50 --warmup 5 --e2e-steps 2
  done ;;
tpnofuse)
  N=$(nvidia-smi -L | wc -l)
  run bench_tp_nofuse_n$N 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --workload tp_decoder_layer --collective symmetric --no-range-fusion --steps 50 --warmup 5 --e2e-steps 2 ;;
tpsweep)
  N=$(nvidia-smi -L | wc -l)
  i=0
  for cfg in "--tp-chunks 4 --reserved-sms 16" "--tp-chunks 4 --reserved-sms 32" "--tp-chunks 8 --reserved-sms 24" ; do
    i=$((i+1))
    run bench_tp_sweep${i}_n$N 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$i bench.py --gpus $N --workload tp_decoder_layer --collective symmetric --steps 50 --warmup 5 --e2e-steps 2 $cfg
  done ;;
esac; done
grep -h '^{' $O/${TAG}_bench_*.log 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d.get('impl','ours'), d['config'].get('workload','')[:60], '| value %.1f' % d['value'], '| ms %.4f' % d['ms_per_step'], '| frac', d.get('roofline',{}).get('frac'), '| e2e %.2f' % d['e2e']['value'], '| clocks', d.get('clocks'))
"
