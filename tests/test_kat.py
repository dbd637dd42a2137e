"""Known-answer tests: the reference's own JUnit vectors, restated in C++ against the host layer
(dcos_commons_b200/host) in tests/kat/*.cpp, and run through the C ABI.

  * not gpu: against the CPU oracle — this is what pins the ORACLE to the reference's golden vectors;
  * gpu:     against the CUDA engine — the same vectors through the product path.
"""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KAT_DIR = os.path.join(ROOT, "tests", "kat")
RUNNER = os.path.join(KAT_DIR, "kat_runner")
ORACLE = os.path.join(ROOT, "oracle", "liboe_oracle.so")
ENGINE = os.path.join(ROOT, "dcos_commons_b200", "liboe_b200.so")


def _build():
    subprocess.check_call(["make", "-C", KAT_DIR, "-s"])
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s"])


def _names():
    _build()
    return subprocess.check_output([RUNNER, "--list"], text=True).split()


NAMES = _names()


def _run(lib, prefix, names, env=None):
    cmd = [RUNNER, "--lib", lib, "--prefix", prefix]
    for n in names:
        cmd += ["--test", n]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, **(env or {})))
    verdict = {}
    for line in p.stdout.splitlines():
        if line.startswith("PASS "):
            verdict[line[5:].strip()] = None
        elif line.startswith("FAIL "):
            name, _, why = line[5:].partition(": ")
            verdict[name] = why
    return p.returncode, verdict, p.stdout + p.stderr


@pytest.fixture(scope="module")
def oracle_verdicts():
    rc, verdict, out = _run(ORACLE, "oeo_", [])
    assert verdict, out
    return verdict


def test_the_suite_restates_enough_vectors():
    assert len(NAMES) >= 170


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_reference_vector(oracle_verdicts, name):
    assert name in oracle_verdicts, f"{name} did not run"
    assert oracle_verdicts[name] is None, oracle_verdicts[name]


@pytest.fixture(scope="module")
def engine_verdicts():
    rc, verdict, out = _run(ENGINE, "oe_", [])
    assert verdict, out
    return verdict


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_engine_matches_reference_vector(engine_verdicts, name):
    assert name in engine_verdicts, f"{name} did not run"
    assert engine_verdicts[name] is None, engine_verdicts[name]


@pytest.mark.gpu
def test_engine_matches_every_vector_with_full_width_offer_tables(engine_verdicts):
    """The host layer sends whole offer tables in their packed form when the library takes it (oe_load_offers_packed); the same
    vectors through the plain oe_load_offers must give the same verdicts."""
    rc, verdict, out = _run(ENGINE, "oe_", [], env={"OE_HOST_WIRE": "plain"})
    assert set(verdict) == set(engine_verdicts), out
    failed = {k: v for k, v in verdict.items() if v is not None}
    assert not failed, failed
