"""Freezes the bitstream md5s of the STOCK encoder (integration/_build/SvtAv1EncApp_stock: the reference built C-only from
an unmodified copy by integration/build_encoder.sh) for the cases of tests/encoder_cases.py. Every case is encoded twice:
a case whose two md5s differ is not reproducible by the reference itself and is refused.

    python tests/golden/make_encoder_md5.py [case ...]
"""
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "svt-av1-psyex_b200", "python"))
sys.path.insert(0, os.path.join(HERE, ".."))

import encoder_cases as ec  # noqa: E402


def main():
    names = sys.argv[1:] or list(ec.CASES)
    try:
        doc = json.load(open(ec.MD5_FILE))
    except Exception:
        doc = {"source": "SvtAv1EncApp built from the unmodified reference (C-only), two runs per case", "md5": {}}
    with tempfile.TemporaryDirectory() as d:
        for name in names:
            a, _ = ec.encode(ec.STOCK, name, d)
            b, _ = ec.encode(ec.STOCK, name, d)
            assert a == b, f"{name}: the stock encoder is not reproducible ({a} vs {b})"
            doc["md5"][name] = a
            print(name, a, flush=True)
            json.dump(doc, open(ec.MD5_FILE, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
