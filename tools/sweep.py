"""Development aid: runs bench.py over kernel shapes (env-selected) and prints one compact line each."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(tag, env, args):
    e = dict(os.environ)
    e.update(env)
    try:
        out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--no-e2e", "--no-cpu-baseline"] + args, env=e,
                             capture_output=True, text=True, timeout=240)
        d = json.loads(out.stdout.strip().splitlines()[-1])
        print(f"{tag:42s} {d['kernel']:12s} ms/step {d['ms_per_step']:9.4f}  {d['value']:.4e} ch-samp/s  frac {d['roofline']['frac']:.3f}  "
              f"clk {d['clocks']['sm_mhz']} {d['clocks']['reasons']}", flush=True)
    except Exception as ex:  # noqa: BLE001
        print(f"{tag:42s} FAILED {ex!r} {out.stderr[-300:] if 'out' in dir() else ''}", flush=True)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "sys"
    if which == "sys":
        big = ["2x16x7x2", "2x16x6x2", "2x8x12x2", "2x32x6x2", "1x8x12x2", "1x16x12x2", "4x16x6x2", "4x32x3x2", "1x16x12x1", "2x16x8x1"]
        for mode in ("fma", "exact"):
            for cfg in big:
                run(f"C65536 {mode} sys {cfg}", {"OLB_BQ_SYS_CFG": cfg}, ["--variant", "3", "--mode", mode, "--steps", "20"])
        small = ["1x16x12x1", "1x8x16x1", "2x16x8x1", "2x16x12x1", "1x8x12x2", "2x16x7x2"]
        for C, F in ((8192, 96000), (4096, 192000)):
            for cfg in small:
                run(f"C{C} fma sys {cfg}", {"OLB_BQ_SYS_CFG": cfg},
                    ["--variant", "3", "--mode", "fma", "--steps", "10", "--channels", str(C), "--block-frames", str(F)])
    if which == "sections":
        for S in (1, 2, 4, 8):
            for v, env in ((2, {}), (3, {"OLB_BQ_SYS_CFG": "2x16x7x2"}), (3, {"OLB_BQ_SYS_CFG": "1x16x12x1"}), (1, {})):
                if v == 3 and S == 1:
                    continue
                run(f"C65536 S{S} variant {v} {env}", env, ["--variant", str(v), "--sections", str(S), "--steps", "20"])
    if which == "modes":
        for mode in ("fast", "fma", "exact"):
            for v, env in ((2, {}), (2, {"OLB_BQ_TMA_CFG": "16x4"}), (3, {"OLB_BQ_SYS_CFG": "2x16x7x2"}), (3, {"OLB_BQ_SYS_CFG": "2x32x6x2"}), (1, {})):
                run(f"C65536 {mode} variant {v} {env}", env, ["--variant", str(v), "--mode", mode, "--steps", "20"])
        for C, F in ((8192, 96000), (4096, 192000)):
            for mode in ("fast", "fma"):
                run(f"C{C} {mode} auto", {}, ["--mode", mode, "--steps", "10", "--channels", str(C), "--block-frames", str(F)])
    if which == "skew":
        shapes = {4096: (192000, ["48x32x4", "48x32x6", "96x16x4", "96x16x6", "48x16x4"]),
                  8192: (96000, ["48x16x4", "48x16x6", "48x32x4", "96x16x4", "48x8x4"]),
                  16384: (48000, ["48x8x4", "48x16x4", "24x8x4"]),
                  32768: (48000, ["24x8x4", "48x8x4", "24x6x4"]),
                  65536: (24000, ["24x4x4", "24x6x4", "24x8x4"])}
        for C, (F, cfgs) in shapes.items():
            for mode in ("fast", "fma", "exact"):
                for cfg in (cfgs if mode == "fast" else cfgs[:1]):
                    run(f"C{C} {mode} skew {cfg}", {"OLB_BQ_SKEW_CFG": cfg},
                        ["--variant", "4", "--mode", mode, "--steps", "10", "--channels", str(C), "--block-frames", str(F), "--no-small"])
            for mode in ("fast", "exact"):
                run(f"C{C} {mode} auto", {}, ["--mode", mode, "--steps", "10", "--channels", str(C), "--block-frames", str(F), "--no-small"])
                run(f"C{C} {mode} tma", {}, ["--variant", "2", "--mode", mode, "--steps", "10", "--channels", str(C), "--block-frames", str(F), "--no-small"])
    if which == "small":
        for C, F in ((4096, 192000), (8192, 96000)):
            for cfg in ("1x32x24x1", "1x64x24x1", "1x32x48x1", "2x64x24x1", "1x32x24x2", "1x32x12x2", "2x32x24x2"):
                for mode in ("fast", "fma"):
                    run(f"C{C} {mode} sys {cfg}", {"OLB_BQ_SYS_CFG": cfg},
                        ["--variant", "3", "--mode", mode, "--steps", "10", "--channels", str(C), "--block-frames", str(F)])
