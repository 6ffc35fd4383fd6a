"""Summarise `nvcc -Xptxas -v` output: registers / spills / smem per kernel.  Usage: make ... 2>&1 | python tools/ptxas_summary.py"""
import re, subprocess, sys
txt = sys.stdin.read()
pat = re.compile(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*?\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\nptxas info\s*: Used (\d+) registers(.*)")
for name, stack, ss, sl, regs, rest in pat.findall(txt):
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    dem = re.sub(r"\(.*", "", dem)
    sm = re.search(r"(\d+) bytes smem", rest)
    print(f"{dem:50s} regs={regs:>3s} stack={stack:>4s} spill_st={ss:>4s} spill_ld={sl:>4s} smem={sm.group(1) if sm else 0}")
