"""Development helper: recompile only inst_n2.cu (the 2-DOF kernels), relink the library and refresh the
content stamp.  Use __graft_entry__.build(force=True) before committing a change that touches shared headers."""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g

os.makedirs(g.OBJ, exist_ok=True)
extra = sys.argv[1:]
obj = os.path.join(g.OBJ, "inst_n2.o")
subprocess.check_call([g.NVCC] + g.NVCC_FLAGS + extra + ["-c", os.path.join(g.CSRC, "inst_n2.cu"), "-o", obj])
objs = [os.path.join(g.OBJ, s.replace(".cu", ".o")) for s in g.SOURCES]
subprocess.check_call([g.NVCC, "-shared", "-o", g.LIB] + objs + ["-lcudart_static", "-lpthread", "-ldl", "-lrt"])
open(g.LIB + ".sha256", "w").write(g._source_hash() + "\n")
print("relinked", g.LIB)
