# A/B of two builds of the library on the GEMM shapes: PENROZ_AB_LIB=<path to the other .so> python tests/probe_ab.py
import os, sys
sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
import penroz._lib as L
if os.environ.get("PENROZ_AB_LIB"):
    L.LIB_PATH = os.path.abspath(os.environ["PENROZ_AB_LIB"])
print("library:", L.LIB_PATH)
exec(open("tests/probe_gemm.py").read())
