"""Mutation fuzzer of the host library's loaders under ASan + UBSan (tools/fuzz/run.sh builds the harness):
    python tools/fuzz/fuzz_stl.py <seed> <cases>
A case may fail to load (exit code 1 with a message); it must never crash or trip a sanitizer."""
import os, random, struct, subprocess, sys, json
random.seed(int(sys.argv[1]) if len(sys.argv) > 1 else 1)
N = int(sys.argv[2]) if len(sys.argv) > 2 else 500
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
WORK = os.environ.get("CRT_FUZZ_DIR", "/tmp/crt_fuzz")
os.makedirs(WORK, exist_ok=True)
meshes = [open(os.path.join(ROOT, "meshes", f), "rb").read() for f in sorted(os.listdir(os.path.join(ROOT, "meshes"))) if f.endswith(".stl") and os.path.getsize(os.path.join(ROOT, "meshes", f)) < 200000][:8]
scene = json.load(open(os.path.join(ROOT, "scenes", "scenetest2.json")))
# find a mesh object or add one
objs = scene["Objects"]
mesh = None
for o in objs:
    if o["type"] == "Mesh": mesh = o
if mesh is None:
    mesh = {"type": "Mesh", "parameters": {"material": objs[0]["parameters"]["material"], "filename": "", "position": [0, 0, 3], "rotation": [0, 0, 0], "scale": 1.0}}
    objs.append(mesh)
mesh["parameters"]["filename"] = os.path.join(WORK, "m.stl")
json.dump(scene, open(os.path.join(WORK, "stl_scene.json"), "w"))
bad = 0
for it in range(N):
    b = bytearray(random.choice(meshes))
    k = random.randrange(7)
    if k == 0: b = b[:random.randrange(len(b) + 1)]
    elif k == 1: b[80:84] = struct.pack("<I", random.choice([0, 1, 0xffffffff, 0x7fffffff, 0x80000000, len(b), (len(b) - 84) // 50 + 1, 0x05555556]))
    elif k == 2:
        for _ in range(random.randint(1, 20)):
            i = random.randrange(84, max(85, len(b) - 4)); b[i:i+4] = struct.pack("<f", random.choice([float("nan"), float("inf"), -float("inf"), 1e38, -1e38, 0.0, 1e-45, 3e12]))
    elif k == 3: b = bytearray()
    elif k == 4: b = b[:84]
    elif k == 5: b = bytearray(b"solid ascii\nfacet normal 0 0 0\n") + b[40:]
    else:
        for _ in range(random.randint(1, 8)): b[random.randrange(len(b))] = random.randrange(256)
    open(os.path.join(WORK, "m.stl"), "wb").write(b)
    r = subprocess.run([os.path.join(WORK, "harness"), "scene", os.path.join(WORK, "stl_scene.json")] + (["norm"] if random.random() < 0.3 else []), cwd=ROOT, capture_output=True, text=True, errors="replace", timeout=120)
    if r.returncode not in (0, 1) or "Sanitizer" in r.stderr or "runtime error" in r.stderr:
        bad += 1
        q = os.path.join(WORK, f"crash_{it}.stl"); open(q, "wb").write(b)
        print("CRASH", it, k, r.returncode, r.stderr.strip().splitlines()[:4], q)
        if bad >= 5: break
print("done", N, "bad", bad)
sys.exit(1 if bad else 0)
