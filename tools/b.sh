#!/bin/bash
# build + load check (run from anywhere)
R=/root/repo
make -C $R/gaussian-processes_b200/csrc -j8 2>&1 | grep -E "error|Error|warning: v" ; \
python -c "
import sys; sys.path.insert(0,'$R/gaussian-processes_b200')
from gpb200 import _lib as G
L=G.load()
missing=[s for s in G.EXPORTS if not hasattr(L,s)]
print('libgpb200 loads; missing exports:', missing)"
