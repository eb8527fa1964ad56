"""Helpers shared by the golden-vector tests: load tests/golden/reference_kats.json and turn the
COO literals into whatever container a backend uses."""
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))


def load_kats():
    with open(os.path.join(HERE, "golden", "reference_kats.json")) as f:
        d = json.load(f)
    return d["fixtures"], d["kats"]


def resolve(fixtures, x):
    return fixtures[x] if isinstance(x, str) else x


def flags(k):
    return dict(replace=k.get("replace", False), comp=k.get("comp", False),
                structure=k.get("structure", False), accum=k.get("accum"))
