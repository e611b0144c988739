#!/usr/bin/env python3
"""Is 1/N of a frame 1/N of the time on ONE GPU?  Renders the whole frame, then each of the N
static shards alone (tile-chunk round robin, the pixels rank k of N would get), and prints the
kernel times: the sum over shards against the whole frame separates what a shard loses on its
own (sparser pixels, fixed cost per launch) from what the devices lose to each other.

    python tools/shard_probe.py scene6 3840 2160 8 -g path -n 64
"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("c-raytracer_b200")
host, device, multigpu, abi = (importlib.import_module(f"c-raytracer_b200.{m}") for m in ("host", "device", "multigpu", "_abi"))


QUICK = bool(os.environ.get("SHARD_PROBE_QUICK"))  # two shards, two frames each, no time lines
REPS = 2 if QUICK else 3


def SHARDS(n):
    return [0, n - 2] if QUICK and n > 2 else list(range(n))


def main():
    scene, w, h, n = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    flags = sys.argv[5:]
    sc = host.Scene(os.path.join(ROOT, "scenes", scene + ".json"))
    dev = device.DeviceScene(sc, 0)
    o = sc.options(w, h, *flags)

    def timed(opts):
        best = None
        for _ in range(REPS):
            st = abi.Stats()
            dev.render(opts, st)
            best = st.kernel_ms if best is None else min(best, st.kernel_ms)
        return best

    whole = timed(o)
    parts = [timed(multigpu.shard_options(sc.options(w, h, *flags), k, n)) for k in SHARDS(n)]
    print(f"{scene} {w}x{h} {' '.join(flags)}: whole frame {whole:.2f} ms; whole/{n} = {whole / n:.2f} ms")

    def timeline(opts, label):
        opts.flags |= abi.CRT_RENDER_COUNT_STATS
        st = abi.Stats()
        dev.render(opts, st)
        print(f"  {label} (instrumented): kernel {st.kernel_ms:.2f} ms; queue dry after {st.queue_ms:.2f} ms, last warp ends at "
              f"{st.last_warp_ms:.2f} ms (tail {st.last_warp_ms - st.queue_ms:.2f} ms), mean warp life {st.warp_busy_ms:.2f} ms")

    if not QUICK:
        timeline(sc.options(w, h, *flags), "whole frame")
        timeline(multigpu.shard_options(sc.options(w, h, *flags), 0, n), f"shard 0 of {n}")
    scale = n / len(parts)
    print(f"  shards alone: {' '.join(f'{p:.2f}' for p in parts)} ms; sum {sum(parts) * scale:.2f} ms ({sum(parts) * scale / whole:.3f} x), max {max(parts):.2f} ms")


if __name__ == "__main__":
    main()
