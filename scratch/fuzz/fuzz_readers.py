"""Mutation fuzz of the on-disk readers (rdyb_exodus_read, rdyb_petsc_vec_read, rdyb_checkpoint_read): truncations and byte flips of the
fixtures must come back as an error code or a valid result, never a crash.  Each trial batch runs in a child process."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
FIX = os.path.join(ROOT, "tests", "golden", "ref_fixtures")

CHILD = os.path.join(ROOT, "tests", "reader_fuzz.py")

def main():
    trials = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    tmp = "/tmp/rdyb_fuzz.bin"
    ck = "/tmp/rdyb_fuzz_src.ck"
    sys.path.insert(0, ROOT)
    import numpy as np
    from rdycore_b200 import io
    io.write_checkpoint(ck, np.arange(90.0).reshape(30, 3), t=1.0, dt=0.1, step=10, nproc=2)
    files = [(os.path.join(FIX, f), "exo") for f in sorted(os.listdir(FIX)) if f.endswith(".exo")]
    files += [(os.path.join(FIX, f), "vec") for f in sorted(os.listdir(FIX)) if f.endswith(".bin")][:4]
    files += [(ck, "ck")]
    bad = 0
    for path, kind in files:
        for seed in range(3):
            r = subprocess.run([sys.executable, CHILD, path, kind, str(seed), str(trials // 3), tmp], capture_output=True, text=True)
            last = r.stdout.strip().splitlines()[-2:] if r.stdout.strip() else []
            if r.returncode != 0:
                bad += 1
                print(f"CRASH rc={r.returncode} {os.path.basename(path)} seed={seed} last={last} err={r.stderr[-300:]}")
            else:
                print(f"ok    {os.path.basename(path)} seed={seed}: {last[-1] if last else ''}")
    print("crashes:", bad)
    return 1 if bad else 0

if __name__ == "__main__":
    sys.exit(main())
