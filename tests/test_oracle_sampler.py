"""Pins for oracle/sampler_ref.py (it used to be checked against nothing).

1. The multinomial draw and the uniform stream against libstdc++ itself: tests/golden/sampler_draws.json holds what
   std::mt19937 + std::discrete_distribution<> -- the two library objects llama.cpp's llama_sample_token uses --
   returned for seeded probability vectors (oracle/draw_ref.cpp, generator tests/golden/make_sampler_golden.py).
   When g++ is on the machine the program is rebuilt and re-run, so a stale fixture cannot hide.
2. The penalty -> top-k -> top-p -> min-p -> temperature chain against small known-answer vectors worked out by
   hand with the published formulas (every expected number below comes with its derivation). The fixed settings the
   reference passes at /root/reference/api.py:59-62 are case F.
"""
import json
import os
import shutil

import numpy as np
import pytest

from conftest import ROOT
from oracle import sampler_ref as sr

GOLD = os.path.join(ROOT, "tests", "golden", "sampler_draws.json")


def test_draw_and_uniform_stream_match_libstdcxx_fixture():
    cases = json.load(open(GOLD))["cases"]
    assert len(cases) >= 10
    for c in cases:
        u = sr.MT19937Uniforms(c["seed"]).take(c["n_draws"])
        p = np.array(c["probs"], np.float32)
        got = [sr.draw(p, float(x)) for x in u]
        assert got == c["draws"], (len(p), c["seed"])


@pytest.mark.skipif(shutil.which("g++") is None, reason="no C++ compiler here; the committed fixture still pins the oracle")
def test_fixture_is_what_libstdcxx_produces_today():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_sampler_golden", os.path.join(ROOT, "tests", "golden", "make_sampler_golden.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    cases = json.load(open(GOLD))["cases"]
    again = m.run_draw_ref(m.build_draw_ref(), cases)
    assert again == [c["draws"] for c in cases]


def test_kat_penalties():
    # history [3, 3, 1, 4]: token 3 twice, tokens 1 and 4 once; repeat 1.1, frequency 0.7, presence 0.8
    #   token 3: 3.0 > 0 -> 3.0 / 1.1 = 2.727273; minus (2 * 0.7 + 0.8) = 2.2          ->  0.527273
    #   token 1: -1.0 <= 0 -> -1.0 * 1.1 = -1.1;  minus (0.7 + 0.8) = 1.5               -> -2.6
    #   token 4: 0.0 <= 0 -> 0.0 * 1.1 = 0;       minus 1.5                             -> -1.5
    out = sr.apply_penalties(np.array([2.0, -1.0, 0.5, 3.0, 0.0], np.float32), [3, 3, 1, 4], 64, 1.1, 0.7, 0.8)
    np.testing.assert_allclose(out, [2.0, -2.6, 0.5, 0.5272727, -1.5], rtol=0, atol=1e-6)
    # only the last `penalty_last_n` tokens count: with last_n = 2 the window is [1, 4]
    out = sr.apply_penalties(np.array([2.0, -1.0, 0.5, 3.0, 0.0], np.float32), [3, 3, 1, 4], 2, 1.1, 0.7, 0.8)
    np.testing.assert_allclose(out, [2.0, -2.6, 0.5, 3.0, -1.5], rtol=0, atol=1e-6)
    # neutral settings leave the logits alone
    out = sr.apply_penalties(np.array([2.0, -1.0], np.float32), [0, 1], 64, 1.0, 0.0, 0.0)
    np.testing.assert_array_equal(out, [2.0, -1.0])


def test_kat_filter_chain_b():
    # logits [1, 4, 2, 3, 0], top-k 3 -> ids [1, 3, 2] with logits [4, 3, 2]
    # softmax: e^0, e^-1, e^-2 = 1, .367879, .135335 (sum 1.503214) -> .665241, .244728, .090031
    # top-p 0.9: .665241 < .9; .665241 + .244728 = .909969 >= .9 -> keep 2
    # min-p 0.05: threshold 4 + ln .05 = 1.004268; 3 >= it -> keep 2
    # temperature 1.2: [3.333333, 2.5] -> e^0, e^-.833333 = 1, .434598 -> .697059, .302941
    ids, p = sr.candidates_after_filters(np.array([1, 4, 2, 3, 0], np.float32), 3, 0.9, 0.05, 1.2)
    assert list(ids) == [1, 3]
    np.testing.assert_allclose(p, [0.6970593, 0.3029407], atol=1e-6)
    # cumulative [.697059, 1]: u = .5 -> first candidate, u = .7 -> second
    assert sr.sample([1, 4, 2, 3, 0], [], 0.5, 1.2, 3, 0.9, 0.05, 1.0) == 1
    assert sr.sample([1, 4, 2, 3, 0], [], 0.7, 1.2, 3, 0.9, 0.05, 1.0) == 3


def test_kat_filter_chain_c_min_p_cuts_after_top_p():
    # logits [5, 1.9, 1.8, -3, .2, 4.9], top-k 40 (all six, order 0 5 1 2 4 3)
    # softmax .500197 .452597 .022533 .020389 .004116 .000168; top-p .99: cumulative .500197 .952793 .975327 .995716 -> keep 4
    # min-p .05: threshold 5 + ln .05 = 2.004268: 4.9 stays, 1.9 goes -> keep 2 (ids 0, 5)
    # temperature .7: [7.142857, 7.0] -> 1, e^-.142857 = .866878 -> .535654, .464346
    ids, p = sr.candidates_after_filters(np.array([5, 1.9, 1.8, -3, 0.2, 4.9], np.float32), 40, 0.99, 0.05, 0.7)
    assert list(ids) == [0, 5]
    np.testing.assert_allclose(p, [0.5356537, 0.4643463], atol=1e-6)


def test_kat_top_p_prefix_rule_and_ties():
    # [2, 1, 0, -1]: softmax .643914 .236883 .087144 .032059. top-p .5 is reached by the first entry alone -> 1 candidate;
    # top-p .65 needs two (.643914 < .65 <= .880797)
    ids, p = sr.candidates_after_filters(np.array([2, 1, 0, -1], np.float32), 4, 0.5, 0.0, 1.0)
    assert list(ids) == [0] and p[0] == 1.0
    ids, _ = sr.candidates_after_filters(np.array([2, 1, 0, -1], np.float32), 4, 0.65, 0.0, 1.0)
    assert list(ids) == [0, 1]
    # equal logits: the lower token id ranks first (the CUDA sampler resolves ties the same way)
    ids, _ = sr.candidates_after_filters(np.array([1, 3, 3, 2, 3], np.float32), 2, 1.0, 0.0, 1.0)
    assert list(ids) == [1, 2]
    # greedy (temperature 0): first maximum
    assert sr.sample([1, 3, 3, 2, 3], [], 0.99, 0.0) == 1


def test_kat_reference_settings_case_f():
    # the sampling the reference always uses (api.py:59-62): temperature 1.2, top_p .9, frequency .7, presence .8,
    # upstream defaults top_k 40, min_p .05, repeat 1.1. logits [2.5, 2.4, .3, -.2, 1.9, 2.45, -1, 0], history [0, 0, 5, 7, 2]
    # penalties: tok 0 (x2): 2.5/1.1 - 2.2 = .072727; tok 5: 2.45/1.1 - 1.5 = .727273; tok 7: 0*1.1 - 1.5 = -1.5;
    #            tok 2: .3/1.1 - 1.5 = -1.227273        -> [.072727, 2.4, -1.227273, -.2, 1.9, .727273, -1, -1.5]
    # order 1 4 5 0 3 6 2 7; softmax .488686 .296403 .091743 .047677 ...; top-p .9: .876832 < .9 <= .924509 -> keep 4
    # min-p: threshold 2.4 + ln .05 = -.595732: .072727 stays -> keep 4
    # / 1.2 -> [2, 1.583333, .606061, .060606] -> .487537 .321404 .120956 .070103 (cumulative .487537 .808941 .929897 1)
    logits = np.array([2.5, 2.4, 0.3, -0.2, 1.9, 2.45, -1.0, 0.0], np.float32)
    hist = [0, 0, 5, 7, 2]
    pen = sr.apply_penalties(logits, hist, 64, 1.1, 0.7, 0.8)
    np.testing.assert_allclose(pen, [0.0727273, 2.4, -1.2272727, -0.2, 1.9, 0.7272727, -1.0, -1.5], atol=1e-6)
    ids, p = sr.candidates_after_filters(pen, 40, 0.9, 0.05, 1.2)
    assert list(ids) == [1, 4, 5, 0]
    np.testing.assert_allclose(p, [0.4875367, 0.3214040, 0.1209559, 0.0701034], atol=1e-6)
    kw = dict(temperature=1.2, top_k=40, top_p=0.9, min_p=0.05, repeat_penalty=1.1, frequency_penalty=0.7, presence_penalty=0.8)
    assert [sr.sample(logits, hist, u, **kw) for u in (0.0, 0.48, 0.49, 0.80, 0.81, 0.92, 0.93, 0.999)] == [1, 1, 4, 4, 5, 5, 0, 0]
