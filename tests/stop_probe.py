"""Developer probe (not a test): the steps of tests/test_gpu_multi.py::test_stop_word_ends_the_other_ranks with progress on stderr."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blt_b200
import oracle
from tests.instances import dense_bilp, reduced_search_input


def say(*a):
    print(*a, file=sys.stderr, flush=True)


A, C = dense_bilp(20, 160, 1)
lat, con, _ = reduced_search_input(A, C)
say("plain")
g = blt_b200.closest(lat, con)
say("plain done", g["rc"], g["found"], g["nodes"])
alone = []
for r in range(2):
    say("alone rank", r)
    alone.append(blt_b200.closest(lat, con, rank=r, world=2, split_depth=3))
    say("  ->", alone[-1]["rc"], alone[-1]["found"], alone[-1]["nodes"])
finder = 0 if alone[0]["found"] else 1
word = (ctypes.c_int * 16)()
addr = ctypes.addressof(word)
say("finder with flag", finder)
first = blt_b200.closest(lat, con, rank=finder, world=2, split_depth=3, stop_flag=addr)
say("  ->", first["rc"], first["found"], first["nodes"], "word", word[0])
say("other with flag set")
second = blt_b200.closest(lat, con, rank=1 - finder, world=2, split_depth=3, stop_flag=addr)
say("  ->", second["rc"], second["found"], second["nodes"], second["stopped"])
