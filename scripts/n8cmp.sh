run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 --steps 1 --warmup 3 2>/dev/null | grep '^{' | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('$2', 'value %.0f ms %.0f passes %d avg %.1f exch %s' % (d['value'], d['ms_per_step'], d['roofline']['launches'], d['roofline']['avg_launch_ms'], d['sharding']))"; }
B200SV_P2P_SWAP=0 run 29513 p2p_off
B200SV_MIN_ROW_BITS=7 run 29514 rows7
B200SV_MIN_ROW_BITS=5 run 29515 rows5
