"""The config surface is the drop-in boundary: ``cmf_b200.config.get_config`` /
``get_schema`` must return, key for key, what the reference's ``config.get_config`` / ``get_schema``
(`config/config.py:39-81`, `config/schemas.py:1-105`) returned for the same dataset and overrides.
The reference's own outputs are stored in every golden fixture by ``oracle/gen_golden.py``."""
import glob
import json
import os

import pytest

from conftest import GOLDEN_DIR, load_golden
from cases import CASES
from cmf_b200.config import get_config, get_schema

NAMES = sorted(n for n in CASES if os.path.exists(os.path.join(GOLDEN_DIR, n + ".pt")))


def _jsonable(o):
    return json.loads(json.dumps(o, default=str))


@pytest.mark.parametrize("name", NAMES)
def test_config_and_schema_equal_the_references(name):
    g = load_golden(name)
    c = CASES[name]
    cfg = {"model": "non-square", "dataset": c["dataset"], **get_config(c["dataset"], "non-square", False), **c["overrides"]}
    ours = _jsonable(cfg)
    assert set(ours) == set(g["config"]), sorted(set(ours) ^ set(g["config"]))
    for k, v in g["config"].items():
        assert ours[k] == v, (k, ours[k], v)
    assert _jsonable(get_schema(cfg)) == g["schema"]


def test_every_fixture_has_a_case():
    have = {os.path.basename(p)[:-3] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.pt"))}
    assert set(CASES) <= have, sorted(set(CASES) - have)
