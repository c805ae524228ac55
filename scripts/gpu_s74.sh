set -x
cd /root/repo
for mb in 128 256; do
timeout 900 python - <<PY 2>&1 | tail -3
import sys, json, torch
sys.path.insert(0, '.')
import bench, whisper_burn_b200 as wb
from whisper_burn_b200.gguf_file import random_whisper_gguf
dev = torch.device('cuda:0')
cfg = wb.WhisperConfig.large_v3()
model = wb.load_whisper_from_gguf(random_whisper_gguf(cfg, seed=0), cfg, dev)
r = bench.leg_transcribe(model, cfg, dev, 0, 1, clips=256, micro=$mb)
print(json.dumps(r))
print('errflag', wb.lib().wq4_device_error_flag(0), 'mem GB', torch.cuda.mem_get_info()[0] / 1e9)
PY
done
