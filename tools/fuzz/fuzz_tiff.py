"""Mutation fuzzer of the host library's loaders under ASan + UBSan (tools/fuzz/run.sh builds the harness):
    python tools/fuzz/fuzz_tiff.py <seed> <cases>
A case may fail to load (exit code 1 with a message); it must never crash or trip a sanitizer."""
import os, random, struct, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
WORK = os.environ.get("CRT_FUZZ_DIR", "/tmp/crt_fuzz")
os.makedirs(WORK, exist_ok=True)
sys.path.insert(0, ROOT)
import _pkg; _pkg.load()
from c_raytracer_b200 import host
import ctypes as C
random.seed(int(sys.argv[1]) if len(sys.argv) > 1 else 1)
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
# seeds: a raw (-f) frame and an 8-bit frame written by the host library
L = host.lib()
w, h = 23, 11
rng = np.random.default_rng(1)
raster = rng.random((h, w, 3), dtype=np.float32); z = rng.random((h, w), dtype=np.float32)
L.rt_host_save.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_void_p, C.c_void_p, C.c_int]
assert L.rt_host_save(os.path.join(WORK, "raw.tif").encode(), w, h, raster.ctypes.data, z.ctypes.data, 1) == 0
assert L.rt_host_save(os.path.join(WORK, "8bit.tif").encode(), w, h, raster.ctypes.data, z.ctypes.data, 0) == 0
seeds = [open(os.path.join(WORK, "raw.tif"), "rb").read(), open(os.path.join(WORK, "8bit.tif"), "rb").read()]
INTS = [0, 1, 2, 3, 0xff, 0xffff, 0x7fffffff, 0x80000000, 0xffffffff, 0xfffffff0, 12, 65000, 273, 279, 256, 257]
bad = 0
for it in range(N):
    b = bytearray(random.choice(seeds))
    k = random.randrange(6)
    if k == 0: b = b[:random.randrange(len(b) + 1)]
    elif k == 1:
        for _ in range(random.randint(1, 6)): b[random.randrange(len(b))] = random.randrange(256)
    elif k in (2, 3):  # overwrite aligned 32-bit / 16-bit words (IFD entries sit at the end or the start) with interesting integers
        for _ in range(random.randint(1, 5)):
            i = random.randrange(0, len(b) - 4)
            if random.random() < 0.5: i = len(b) - 4 - random.randrange(0, min(len(b) - 4, 260))
            if random.random() < 0.5: b[i:i+4] = struct.pack("<I", random.choice(INTS))
            else: b[i:i+2] = struct.pack("<H", random.choice(INTS) & 0xffff)
    elif k == 4: b[4:8] = struct.pack("<I", random.choice(INTS + [len(b) - 2, len(b), len(b) + 100, 8]))
    else: b = b + bytes(random.randrange(256) for _ in range(random.randint(1, 64)))
    open(os.path.join(WORK, "m.tif"), "wb").write(b)
    r = subprocess.run([os.path.join(WORK, "harness"), "tiff", os.path.join(WORK, "m.tif")], capture_output=True, text=True, errors="replace", timeout=60)
    if r.returncode not in (0, 1) or "Sanitizer" in r.stderr or "runtime error" in r.stderr:
        bad += 1
        q = os.path.join(WORK, f"crash_{it}.tif"); open(q, "wb").write(b)
        print("CRASH", it, k, r.returncode, r.stderr.strip().splitlines()[:5], q)
        if bad >= 5: break
print("done", N, "bad", bad)
sys.exit(1 if bad else 0)
