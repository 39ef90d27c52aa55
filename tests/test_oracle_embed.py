"""CPU tests for the embedder row: the oracle against the committed golden vectors (tokens from the
reference's own tokeniser, embeddings from torch.nn.GRU on the reference's weights - tests/golden/make_embedder_golden.py),
the host-side pieces of the product (tokeniser, IR reader) and the fail-loudly rule."""
import os

import numpy as np
import pytest

import deepreadmapper_b200 as drm
from oracle import embed_oracle as EO

import embed_util as U

REF_MODEL = "/root/reference/models/finetuned_sgn33-new-a-Apr6"


@pytest.fixture(scope="module")
def gold():
    return U.load_golden()


def test_token_ids_follow_the_reference_table():
    # spot values of src/utils/tok2index.cpp and the size of the vocabulary they imply
    assert [EO.token_id(s) for s in (0, 15, 16, 17, 31, 32, 35, 36, 95)] == [7542, 7557, 7558, 7563, 7633, 7559, 7562, 7564, 7637]
    assert len({EO.token_id(s) for s in range(96)}) == 96


def test_oracle_tokeniser_matches_reference_tokens(gold):
    assert np.array_equal(EO.tokenise_batch(gold["tok_seqs"]), gold["tok_all"])


def test_product_tokeniser_matches_reference_tokens(gold):
    assert np.array_equal(drm.tokenize(gold["tok_seqs"]), gold["tok_all"])


def test_product_tokeniser_matches_oracle_on_ragged_and_undefined_input():
    g = np.random.default_rng(5)
    seqs = [b"", b"A", b"AC", b"<A>", b"<>", b"ACGTNACGT", b"<ACGNNT>", b"NN", b"acgtXacgt>" * 20,
            b"<" + bytes(g.choice(np.frombuffer(b"ACGTN", np.uint8), 200)) + b">"]
    for ln in (2, 3, 122, 123, 124, 125, 300):
        seqs.append(bytes(g.choice(np.frombuffer(b"ACGTacgt", np.uint8), ln)))
    for T in (123, 16, 128):
        assert np.array_equal(drm.tokenize(seqs, max_len=T), EO.tokenise_batch(seqs, T))
    # the decorated form the reference feeds: first token = hash('<', '<', b0) = the "t?>" row of its table
    assert drm.tokenize([b"<ACGT>"])[0, :6].tolist() == [7618, 7543, 7566, 7592, 7613, 7562]


def test_oracle_network_matches_torch_golden(gold):
    got = EO.embed_tokens(gold["tokens"], gold["weights"])
    assert got.shape == gold["embeddings"].shape == (len(gold["seqs"]), 128)
    assert float(np.abs(got - gold["embeddings"]).max()) < 2e-5
    # and end to end from the sequences
    got2 = EO.embed(gold["seqs"], gold["weights"])
    assert np.array_equal(got, got2)


def test_ir_reader_on_a_written_model(tmp_path):
    w = EO.random_weights(3)
    for f32 in ((), ("B1", "R2", "emb.weight")):
        xml = U.write_ir(str(tmp_path / f"m{len(f32)}"), w, f32=f32)
        vocab, counts, sums = drm.embedder_ir_info(xml, xml[:-4] + ".bin")
        assert vocab == EO.VOCAB
        assert counts.tolist() == [w[k].size for k in ("emb", "W1", "R1", "B1", "W2", "R2", "B2")]
        want = [float(w[k].astype(np.float64).sum()) for k in ("emb", "W1", "R1", "B1", "W2", "R2", "B2")]
        assert np.allclose(sums, want, rtol=0, atol=1e-6)


def test_ir_reader_rejects_other_architectures(tmp_path):
    w = EO.random_weights(4)
    xml = U.write_ir(str(tmp_path / "m"), w)
    text = open(xml).read()
    open(xml, "w").write(text.replace('linear_before_reset="true"', 'linear_before_reset="false"'))
    with pytest.raises(drm.DrmError) as e:
        drm.embedder_ir_info(xml, xml[:-4] + ".bin")
    assert e.value.code == -4
    open(xml, "w").write(text)
    open(xml[:-4] + ".bin", "wb").write(open(xml[:-4] + ".bin", "rb").read()[:1000])  # truncated weights
    with pytest.raises(drm.DrmError) as e:
        drm.embedder_ir_info(xml, xml[:-4] + ".bin")
    assert e.value.code == -2
    with pytest.raises(drm.DrmError) as e:
        drm.embedder_ir_info(str(tmp_path / "missing.xml"), str(tmp_path / "missing.bin"))
    assert e.value.code == -2


def test_ir_reader_survives_malformed_numbers(tmp_path):
    """Attributes that std::stoull / stoll / stoi refuse, sizes that overflow, negative and zero dimensions: an error code,
    never an exception through the C ABI (which would end the process under ctypes)."""
    import re
    w = EO.random_weights(5)
    xml = U.write_ir(str(tmp_path / "m"), w)
    text = open(xml).read()
    m_off, m_shape = re.search(r'offset="(\d+)"', text), re.search(r'shape="7638, 64"', text)
    assert m_off and m_shape
    cases = [text.replace(m_off.group(0), 'offset="abc"', 1),
             text.replace(m_off.group(0), 'offset="99999999999999999999999999"', 1),
             text.replace(m_off.group(0), 'offset="18446744073709551615"', 1),
             text.replace(m_shape.group(0), 'shape="x, 64"'),
             text.replace(m_shape.group(0), 'shape="-1, 64"'),
             text.replace(m_shape.group(0), 'shape="0, 64"'),
             text.replace(m_shape.group(0), 'shape="4611686018427387904, 64"'),
             re.sub(r'<layer id="(\d+)"', '<layer id="q\\1"', text, count=1),
             re.sub(r'from-layer="(\d+)"', 'from-layer="?"', text, count=1)]
    for i, bad in enumerate(cases):
        assert bad != text, i
        open(xml, "w").write(bad)
        with pytest.raises(drm.DrmError) as e:
            drm.embedder_ir_info(xml, xml[:-4] + ".bin")
        assert e.value.code in (-2, -4), (i, e.value.code, str(e.value))


@pytest.mark.skipif(not os.path.exists(REF_MODEL + ".xml"), reason="needs /root/reference (build container only)")
def test_ir_reader_on_the_reference_model(gold):
    vocab, counts, sums = drm.embedder_ir_info(REF_MODEL + ".xml", REF_MODEL + ".bin")
    w = EO.load_ir(REF_MODEL + ".xml", REF_MODEL + ".bin")
    assert vocab == 7638
    want = [float(w[k].astype(np.float64).sum()) for k in ("emb", "W1", "R1", "B1", "W2", "R2", "B2")]
    assert np.allclose(sums, want, rtol=0, atol=1e-6)
    # the committed fixture is a subset of that model
    for k in ("W1", "R1", "B1", "W2", "R2", "B2"):
        assert np.array_equal(w[k], gold["weights"][k])


def test_embedder_fails_loudly_without_gpu(gold):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(drm.DrmError) as e:
        drm.Vectorizer(weights=gold["weights"])
    assert e.value.code == -3 and "no CPU fallback" in str(e.value)
