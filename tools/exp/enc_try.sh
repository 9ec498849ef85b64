python - <<'PY'
import sys,tempfile
sys.path.insert(0,'svt-av1-psyex_b200/python'); sys.path.insert(0,'tests')
import encoder_cases as ec
from me_b200 import abi
try:
    print(ec.encode(ec.GPU, "640x360_m8_scene", "/tmp", env={"SVT_ME_GPU_LIB": abi.GPU_LIB})[0])
except Exception as e:
    print(str(e)[-600:])
PY
