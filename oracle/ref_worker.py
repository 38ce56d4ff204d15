"""Child-process entry of oracle.ref_py.run_isolated — TEST INFRASTRUCTURE ONLY."""
import os
import pickle
import sys


def main():
    fin, fout = sys.argv[1], sys.argv[2]
    from oracle import ref_py
    with open(fin, "rb") as f:
        spec = pickle.load(f)
    out = ref_py.run_ops(spec)
    with open(fout + ".tmp", "wb") as f:
        pickle.dump(out, f, protocol=4)
    os.replace(fout + ".tmp", fout)
    sys.stdout.flush()
    os._exit(0)  # skip interpreter teardown: the reference may have corrupted the heap


if __name__ == "__main__":
    main()
