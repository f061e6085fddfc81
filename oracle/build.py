"""Builds the oracle's C restatement (oracle/_ref/liboracle.so).  Test infrastructure only.
The reference itself is pure Python (zero native code), so there is no reference
C/C++ to compile into oracle/_ref; the directory only holds this restatement's binary."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, '_ref', 'liboracle.so')


def build(force=False):
    src = os.path.join(HERE, 'nms_greedy.c')
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= os.path.getmtime(src):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    subprocess.check_call(['gcc', '-O2', '-ffp-contract=off', '-fno-fast-math', '-shared', '-fPIC', src, '-o', OUT])
    return OUT


if __name__ == '__main__':
    print(build(force=True))
