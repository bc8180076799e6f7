"""Register / stack / spill table from an `nvcc -Xptxas -v` log (stdin or file)."""
import re
import subprocess
import sys

t = open(sys.argv[1]).read() if len(sys.argv) > 1 else sys.stdin.read()
for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*?\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores.*?\n.*?Used (\d+) registers", t):
    name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
    name = re.sub(r"\(.*", "", name)
    print(f"{m.group(4):>4} regs {m.group(2):>4} stack {m.group(3):>3} spill  {name}")
