"""A/B of two builds of the library on one box: python scratch/ab_lib.py <libgomoku_b200_variant.so> <bench.py arguments>.
The variant is loaded first; bench.py's own engine.load_library() then returns it."""
import os, runpy, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path[:0] = [os.path.join(ROOT, "alphazero-gomoku_b200"), ROOT]
import engine
engine.load_library(os.path.abspath(sys.argv[1]))
sys.argv = [os.path.join(ROOT, "bench.py")] + sys.argv[2:]
runpy.run_path(sys.argv[0], run_name="__main__")
