"""oracle/patch_beta.py -- TEST INFRASTRUCTURE.  Writes a scratch copy of the reference's eri_sparse_DF.cpp in which
the cublasDsyrk of eval_via_cholesky_cuda (cpp:684-694) is called with beta = 0 instead of beta = 1 on an
uninitialised buffer, for the "beta-fixed" GPU bar.  usage: patch_beta.py <reference.cpp> <scratch.cpp>"""
import re
import sys

src = open(sys.argv[1]).read()
m = re.search(r"cublasDsyrk\(.*?\)\);", src, flags=re.S)
assert m, "cublasDsyrk call not found"
call = m.group(0)
parts = call.split("&alpha")
assert len(parts) == 3, "expected alpha and beta both passed as &alpha"
fixed = parts[0] + "&alpha" + parts[1] + "&qemb_beta_zero" + parts[2]
src = src.replace(call, fixed, 1)
anchor = "const double alpha = 1.0;"
assert src.count(anchor) == 1
src = src.replace(anchor, anchor + "\n    const double qemb_beta_zero = 0.0;", 1)
open(sys.argv[2], "w").write(src)
