"""Checks tests/golden/reference_kats.json against the oracle and re-derives its derived fields (the varint / string
bytes Go's binary.PutVarint produces for the values the reference's tests use).  The reference is Go and cannot be
executed in this image, so the fixture is a transcription of the assertions in the reference's *_test.go files; this
script is the committed generator for everything in it that is computed rather than transcribed."""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import binding as oracle  # noqa: E402

oracle.build()
path = ROOT / "tests" / "golden" / "reference_kats.json"
kats = json.loads(path.read_text())
for v in kats["varint"]:
    b = oracle.put_varint(v["value"])
    v["bytes_hex"], v["n_bytes"] = b.hex(), len(b)
for s in kats["string"]:
    raw = s["value"].encode()
    b = oracle.put_varint(len(raw)) + raw  # encode.go:77-88: the length goes through binary.PutVarint (zig-zag)
    s["bytes_hex"], s["n_bytes"] = b.hex(), len(b)
path.write_text(json.dumps(kats, indent=1) + "\n")
print("golden fixture refreshed:", path)
