#!/usr/bin/env python
"""Print the ctypes declarations of the ABI structs exactly as speech_recognition_tools_b200/_lib.py binds them.
INTEGRATION.md embeds this output between its GENERATED markers; tests/test_host_logic.py checks that the document,
_lib.py and include/fdlp_b200.h agree field by field.

    python tools/gen_binding_stub.py            # print
    python tools/gen_binding_stub.py --write    # refresh INTEGRATION.md in place
"""
import ctypes as C
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speech_recognition_tools_b200 import _lib as L  # noqa: E402

BEGIN, END = "<!-- BEGIN GENERATED ctypes structs (tools/gen_binding_stub.py) -->", "<!-- END GENERATED -->"
C_NAMES = {"FdlpParams": "fdlp_params", "FdlpBatch": "fdlp_batch"}


FIXED = {"c_int": "c_int32", "c_uint": "c_uint32", "c_long": "c_int64", "c_ulong": "c_uint64"}


def ctype_name(t):
    if isinstance(t, type) and issubclass(t, C.Array):
        return "C.%s * %d" % (FIXED.get(t._type_.__name__, t._type_.__name__), t._length_)
    if isinstance(t, type) and issubclass(t, C._Pointer):
        return "C.POINTER(C.%s)" % FIXED.get(t._type_.__name__, t._type_.__name__)
    return "C." + FIXED.get(t.__name__, t.__name__)


def stub():
    out = ["```python", "import ctypes as C", ""]
    for py, cn in C_NAMES.items():
        cls = getattr(L, py)
        out.append("class %s(C.Structure):      # include/fdlp_b200.h: %s, %d bytes" % (cn, cn, C.sizeof(cls)))
        out.append("    _fields_ = [")
        for name, t in cls._fields_:
            out.append("        (%r, %s)," % (name, ctype_name(t)))
        out.append("    ]")
        out.append("")
    out.append("```")
    return "\n".join(out)


def main():
    text = stub()
    if "--write" in sys.argv:
        path = os.path.join(ROOT, "INTEGRATION.md")
        doc = open(path).read()
        pat = re.compile(re.escape(BEGIN) + ".*?" + re.escape(END), re.S)
        assert pat.search(doc), "markers missing in INTEGRATION.md"
        open(path, "w").write(pat.sub(lambda m: BEGIN + "\n" + text + "\n" + END, doc))
    else:
        print(text)


if __name__ == "__main__":
    main()
