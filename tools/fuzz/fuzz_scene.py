"""Mutation fuzzer of the host library's loaders under ASan + UBSan (tools/fuzz/run.sh builds the harness):
    python tools/fuzz/fuzz_scene.py <seed> <cases>
A case may fail to load (exit code 1 with a message); it must never crash or trip a sanitizer."""
import os, random, subprocess, sys, re
random.seed(int(sys.argv[1]) if len(sys.argv) > 1 else 1)
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
WORK = os.environ.get("CRT_FUZZ_DIR", "/tmp/crt_fuzz")
os.makedirs(WORK, exist_ok=True)
scenes = [open(os.path.join(ROOT, "scenes", f)).read() for f in ("scene1.json", "scene2.json", "scene3.json", "scenetest.json", "scenetest2.json", "scenetest3.json", "scene5_small.json")]
EXTREME = ["1e999", "-1e999", "0", "-0.0", "1e-320", "nan", "NaN", "1e38", "-1e38", "4294967296", "-1", "2147483648", "99999999999999999999", "0.0000001", "[]", "{}", "null", "true", '"x"', "1e12", "3e12"]
def mutate(t):
    k = random.randrange(9)
    if k == 0:  # flip bytes
        b = bytearray(t.encode())
        for _ in range(random.randint(1, 4)):
            b[random.randrange(len(b))] = random.randrange(256)
        return b.decode("latin1")
    if k == 1:  # delete a span
        i = random.randrange(len(t)); j = min(len(t), i + random.randint(1, 200))
        return t[:i] + t[j:]
    if k == 2:  # duplicate a span
        i = random.randrange(len(t)); j = min(len(t), i + random.randint(1, 400))
        return t[:j] + t[i:j] * random.randint(1, 3) + t[j:]
    if k == 3:  # truncate
        return t[:random.randrange(len(t))]
    if k in (4, 5, 6):  # replace a number by an extreme value
        nums = list(re.finditer(r"-?\d+\.?\d*(?:[eE][-+]?\d+)?", t))
        if not nums: return t
        out = t
        for m in random.sample(nums, min(len(nums), random.randint(1, 3))):
            out = out[:m.start()] + random.choice(EXTREME).ljust(m.end() - m.start())[:max(m.end() - m.start(), 1)] + out[m.end():] if random.random() < 0.3 else out[:m.start()] + random.choice(EXTREME) + out[m.end():]
            break
        return out
    if k == 7:  # deep nesting
        i = random.randrange(len(t))
        return t[:i] + "[" * random.randint(100, 5000) + t[i:]
    # swap a key / string
    strs = list(re.finditer(r'"[^"]*"', t))
    if not strs: return t
    m = random.choice(strs)
    return t[:m.start()] + random.choice(['""', '"type"', '"Mesh"', '"Sphere"', '"Triangle"', '"Plane"', '"material"', '"' + "A" * 5000 + '"', '"meshes/nonexistent.stl"', '"meshes/cube.stl"', '"\\u0000"', '"\\ud800"', '"\\']) + t[m.end():]
bad = 0
for it in range(N):
    t = mutate(random.choice(scenes))
    if random.random() < 0.3: t = mutate(t)
    p = os.path.join(WORK, "m.json")
    open(p, "w", encoding="latin1", errors="replace").write(t)
    r = subprocess.run([os.path.join(WORK, "harness"), "scene", p] + (["norm"] if random.random() < 0.2 else []), cwd=ROOT, capture_output=True, text=True, errors="replace", timeout=60)
    if r.returncode not in (0, 1) or "Sanitizer" in r.stderr or "runtime error" in r.stderr:
        bad += 1
        q = os.path.join(WORK, f"crash_{it}.json"); open(q, "w", encoding="latin1", errors="replace").write(t)
        print("CRASH", it, r.returncode, r.stderr.strip().splitlines()[:3], q)
        if bad >= 5: break
print("done", N, "bad", bad)
sys.exit(1 if bad else 0)
