for cfg in "32 3" "12 3" "96 3" "32 2" "32 5" "8 2"; do set -- $cfg; DEFL_CHUNK_COUNT=$1 DEFL_PIPE_SLOTS=$2 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus 8 --steps 3 --warmup 3 --no-extra --no-cpu 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('chunks $1 slots $2', round(j['e2e']['value'],1), round(j['e2e']['copy_ceiling'],1), round(j['e2e']['of_copy_ceiling'],3))"; done
