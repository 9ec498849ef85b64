cd /tmp && python - <<'PY' 2>&1 | tail -30
import sys, os
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo/svt-av1-psyex_b200/python")
import encoder_cases as ec
from me_b200 import abi
try:
    md5, log = ec.encode(ec.GPU, "320x192_m8", "/tmp", env={"SVT_ME_GPU_LIB": abi.GPU_LIB, "SVT_ME_GPU_LOG_LISTS": "1"})
    print(md5)
    print("\n".join(l for l in log.splitlines() if "tf pair" in l or "svt_me_gpu_hook: " in l and "lists" not in l)[:3000])
except RuntimeError as e:
    print(str(e)[:300]); print(str(e)[-600:])
PY
