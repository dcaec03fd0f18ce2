"""Helper (not a test module): mutate one fixture file n times — truncation, header byte flips, 32-bit words set to extreme or to small in-range values —
and pass each mutant to the reader.  A reader must answer with an error (RuntimeError from the C-ABI code) or a valid result; a
crash ends this process with a signal, which the caller (tests/test_io_formats.py, scratch/fuzz/fuzz_readers.py) reports.
Usage: python tests/reader_fuzz.py <fixture> <exo|vec|ck> <seed> <n> <tmpfile>"""
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rdycore_b200 import io  # noqa: E402


def main(src, kind, seed, n, tmp):
    data = open(src, "rb").read()
    rng = random.Random(seed)
    ok = err = 0
    for t in range(n):
        b = bytearray(data)
        mode = rng.randrange(4)
        if mode == 0:
            b = b[:rng.randrange(len(b))]
        elif mode == 1:
            for _ in range(rng.randrange(1, 6)):
                b[rng.randrange(min(len(b), 4096))] = rng.randrange(256)  # header bytes
        elif mode == 2:
            for _ in range(rng.randrange(1, 6)):
                i = rng.randrange(len(b) - 4)
                b[i:i + 4] = rng.choice([b"\xff\xff\xff\xff", b"\x7f\xff\xff\xff", b"\x00\x00\x00\x00", b"\x80\x00\x00\x00"])
        else:  # small valid-looking integers on aligned words: node / element / side ids that stay in range but break the topology
            for _ in range(rng.randrange(1, 40)):
                i = rng.randrange((len(b) - 4) // 4) * 4
                b[i:i + 4] = rng.randrange(1, 60).to_bytes(4, "big")
        with open(tmp, "wb") as f:
            f.write(bytes(b))
        print(t, mode, flush=True)
        try:
            out = {"exo": io.read_exodus, "vec": io.read_petsc_vec, "ck": io.read_checkpoint}[kind](tmp)
            if kind == "exo":  # a file that parses may still hold a broken topology: the mesh builder must refuse it, not crash
                mesh = io.mesh_from_exodus(out)
                io.boundaries_from_side_sets(mesh, out)
            ok += 1
        except (RuntimeError, MemoryError, ValueError):
            err += 1
    print("done", ok, err)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5])
