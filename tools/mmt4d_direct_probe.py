"""Debug probe: one tile-direct mmt4d call per process (a fault kills the context)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from iree_b200 import hal as H

t, M, K, N, M0, N0, acc = sys.argv[1], *(int(x) for x in sys.argv[2:8])
K0 = 16 if t == "i8" else 8
dev = H.Device(0); ops = H.Ops(dev)
lhs = oracle.generate_random_matrix(t, (M, K), 1)
rhs = oracle.generate_random_matrix(t, (K, N), 2)
l4 = oracle.pack(lhs, M0, K0); r4 = oracle.pack(np.ascontiguousarray(rhs.T), N0, K0)
M1, K1, N1 = l4.shape[0], l4.shape[1], r4.shape[0]
dt = np.int32 if t == "i8" else np.float32
a = oracle.generate_random_matrix("i32" if t == "i8" else "f32", (M1, N1, M0, N0), 3) if acc else None
out = a.copy() if acc else np.full((M1, N1, M0, N0), 77, dtype=dt)
before = ops.dispatch_count
ops.mmt4d_host(t, M1, N1, K1, M0, N0, K0, l4.ctypes.data, r4.ctypes.data, out.ctypes.data,
               flags=H.MATMUL_ACCUMULATE if acc else 0)
want = oracle.mmt4d(t, l4, r4, acc=a)
print(sys.argv[1:], "dispatches", ops.dispatch_count - before, "equal", np.array_equal(out, want),
      "mismatches", int((out != want).sum()), flush=True)
