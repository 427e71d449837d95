"""The default transpose kernel against the same kernel without its turn through shared memory (la_transpose_set_mode(3):
identical global reads and writes, wrong bytes inside each tile): what the memory system gives the access pattern."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
exec(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "tr_experiments.py")).read().split("lib.la_transpose_set_variant(0); line(")[0])
lib.la_transpose_set_variant(0)
lib.la_transpose_set_mode(0); line("LDG/STG 64x64, the transpose")
lib.la_transpose_set_mode(3); line("LDG/STG 64x64, no turn (traffic only)")
lib.la_transpose_set_mode(0)
