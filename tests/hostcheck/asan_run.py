"""The host loops over the transfer's per-item routines (the source the CUDA kernels run) under
AddressSanitizer: compute-sanitizer is closed on the GPU pool, so out-of-bounds accesses of the
shared routines are looked for here -- the recorded reference case, the k-NN point sets
(random, lattice, flat, degenerate) and the tiny meshes.

  g++ -O1 -g -fsanitize=address -fno-omit-frame-pointer -ffp-contract=off -fopenmp -fPIC -shared -std=c++17 \
      -o /tmp/hc_asan/libtransfer_hostcheck.so tests/hostcheck/transfer_hostcheck.cpp
  LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0 OMP_NUM_THREADS=4 \
      python tests/hostcheck/asan_run.py /tmp/hc_asan/libtransfer_hostcheck.so
"""
import ctypes as C
import os
import pathlib
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import test_transfer as T  # noqa: E402


def main():
    lib = sys.argv[1]

    def hc():
        L = C.CDLL(lib); V = C.c_void_p
        L.hostcheck_transfer_cells.argtypes = [V, C.c_int, C.c_int, V, V, V, V]
        L.hostcheck_transfer_nodes.argtypes = [V, V, V, C.c_int, V, V, V, V]
        L.hostcheck_knn.argtypes = [C.c_int, V, C.c_int, V, C.c_int, V, V]
        return L
    T.hostcheck = hc
    T.test_per_item_routines_reproduce_the_reference_transfer()
    for k in ("random", "lattice", "flat"):
        T.test_grid_knn_equals_brute_force(k)
    for k in ("one-point", "coincident", "collinear", "two-cells-far-apart"):
        T.test_grid_knn_on_degenerate_point_sets(k)
    for r in (6e3, 3e3):
        try:
            T.test_per_item_routines_on_tiny_meshes(r, pathlib.Path(tempfile.mkdtemp()))
        except BaseException as e:      # pytest.skip when oracle/_ref is absent
            if "not built" not in str(e):
                raise
    print("ASAN RUN COMPLETE: no report")


if __name__ == "__main__":
    main()
