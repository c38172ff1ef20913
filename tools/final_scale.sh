set -x
for n in 1 2 4 8; do
  if [ $n = 1 ]; then python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/r01g_scale_n1.json 2> gpurun_out/r01g_scale_n1.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps 100 --warmup 5 > gpurun_out/r01g_scale_n$n.json 2> gpurun_out/r01g_scale_n$n.err; fi
  tail -c 300 gpurun_out/r01g_scale_n$n.json | head -c 10 > /dev/null
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29700 tools/gan_dp_check.py > gpurun_out/r01g_gan_dp_n8.log 2>&1
grep "rank 0\|weak" gpurun_out/r01g_gan_dp_n8.log
python - <<'PY'
import json
for n in (1,2,4,8):
    try:
        d=json.loads(open(f'gpurun_out/r01g_scale_n{n}.json').read().strip().splitlines()[-1])
        print(n, d['value'], d['ms_per_step'], d['e2e']['value'], d.get('clocks',{}).get('reasons'))
    except Exception as e:
        print(n, 'ERR', e)
PY
