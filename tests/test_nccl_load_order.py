"""libhnswb200 dlopens NCCL at first multi-GPU use; a Python process that imports PyTorch afterwards must still get a
libnccl that has every symbol libtorch_cuda needs (the dynamic linker keys loaded objects by SONAME).  CPU only:
ncclGetUniqueId needs no device."""
import subprocess
import sys
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CODE = """
import ctypes as C, sys
sys.path.insert(0, %r)
from hnsw_itu_b200 import _lib
L = _lib.lib()
ident = (C.c_uint8 * 128)()
rc = L.hb_comm_unique_id(ident)
assert rc == 0, (rc, L.hb_last_error())
assert any(bytes(ident)), "empty unique id"
import torch   # must not fail on an older libnccl loaded first
print("ok")
""" % ROOT


def test_importing_torch_after_the_library_loaded_nccl():
    env = {k: v for k, v in os.environ.items() if k != "HB_NCCL_LIB"}
    r = subprocess.run([sys.executable, "-c", CODE], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stderr[-2000:]
