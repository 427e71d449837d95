import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, ".")
exec(open("tools/tr_experiments.py").read().split("lib.la_transpose_set_variant(0); line(")[0])
lib.la_transpose_set_variant(0)
for band in (-1, 1, 2, 4, 8, 16, 32, 64):
    lib.la_transpose_set_band(band); line("LDG/STG 64x64 band %d" % band)
lib.la_transpose_set_band(-1)
