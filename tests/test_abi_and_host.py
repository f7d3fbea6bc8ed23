"""CPU-side checks: the C-ABI library loads and exports every symbol include/retr_b200.h declares
(no compute calls), and the host-side mirrors of the reference's helper functions behave like
the reference (eval_utils/decode.py, engine.py, models/utils.py)."""
import os
import re

import pytest
import torch

from conftest import ROOT, build_product_model
from oracle import retr_oracle as O


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "retr_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(retr_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    import __graft_entry__
    __graft_entry__.build()
    from retr_b200 import native
    lib = native.lib()
    names = declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/retr_b200.h but not exported"
        assert n in native.SIGNATURES, f"{n} has no ctypes signature"
    assert sorted(native.SIGNATURES) == names
    assert lib.retr_abi_version() == 1
    assert isinstance(lib.retr_last_error(), bytes)


def test_no_cuda_means_loud_failure():
    """The product path has no CPU fallback: a CPU-resident model refuses to run."""
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    model, _, _ = build_product_model("Caption")
    from retr_b200 import synthetic
    data = synthetic.make_batch(1, train=True)
    with pytest.raises(RuntimeError, match="CUDA"):
        model.eval()(*synthetic.pack(data, "Caption"), data["caps"][:, :-1], data["cap_masks"][:, :-1])
    from retr_b200.eval_utils import decode
    with pytest.raises(RuntimeError, match="CUDA"):
        decode.greedy(synthetic.pack(data, "Caption"), model, max_len=128, device="cpu", bos_token=101, eos_token=102)


def test_product_never_imports_oracle():
    """oracle/ is test infrastructure: nothing under retr_b200/ may import it."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "retr_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, os.path.join(dirpath, f)


def test_create_caption_and_mask_and_prune():
    from retr_b200.eval_utils import decode
    cap, msk = decode.create_caption_and_mask(101, 128, 3)
    rcap, rmsk = O.create_caption_and_mask(101, 128, 3)
    assert torch.equal(cap, rcap) and torch.equal(msk, rmsk)
    rows = [[101, 5, 6, 102, 7, 0], [101, 9, 0, 0, 0, 0], [101, 102, 5, 5, 5, 5], [101, 7, 7, 7, 7, 7]]
    for clean in (True, False):
        assert decode.prune_cap_ids(rows, clean, 0, 101, 102) == O.prune_cap_ids(rows, clean, 0, 101, 102)
    assert decode.prune_cap_ids(rows, True, 0, 101, 102) == [[5, 6], [9], [], [7, 7, 7, 7, 7]]


def test_literal_greedy_with_foreign_callable():
    """A callable that is not a retr_b200 model gets the reference's literal loop, including the
    early return before the last EOS is written (eval_utils/decode.py:73-79)."""
    from retr_b200.eval_utils import decode
    from retr_b200.models.utils import NestedTensor
    V = 50
    script = {0: [7, 9, 102], 1: [8, 102, 4]}        # row -> token per step

    def stub(samples, caption, cap_mask):
        B, T = caption.shape
        out = torch.zeros(B, T, V + 60)
        filled = int((~cap_mask[0]).sum())            # positions written so far
        for b in range(B):
            tok = script[b][min(filled - 1, 2)]
            out[b, filled - 1, tok] = 1.0
        return out

    s = (NestedTensor(torch.zeros(2, 3, 4, 4), torch.zeros(2, 4, 4, dtype=torch.bool)),)
    ids = decode.greedy(s, stub, max_len=8, device="cpu", bos_token=101, eos_token=102)
    # row 1 finishes at step 1 (EOS written), row 0 at step 2 -> return without writing step 2
    assert ids.tolist() == [[101, 7, 9, 0, 0, 0, 0, 0], [101, 8, 102, 0, 0, 0, 0, 0]]


def test_pack_encoder_inputs_shapes():
    from retr_b200 import engine
    t, tm = torch.zeros(2, 3, 8, 8), torch.zeros(2, 8, 8, dtype=torch.bool)
    loc = torch.zeros(2, 5)
    assert len(engine.pack_encoder_inputs((t, tm), False, False)) == 1
    out = engine.pack_encoder_inputs((t, tm, loc), False, True)
    assert len(out) == 2 and out[1].shape == (2, 5)
    out = engine.pack_encoder_inputs((t, tm, t, tm), True, False)
    assert len(out) == 2 and out[1].shape == tm.shape            # NestedTensor.shape is the mask shape
    out = engine.pack_encoder_inputs((t, tm, t, tm, loc), True, True)
    assert len(out) == 3 and out[0].shape[0] == 2


def test_build_model_flag_matrix():
    from retr_b200.configuration import Config
    from retr_b200.models import caption
    with pytest.raises(NotImplementedError):
        caption.build_model(Config(use_global_features=True, use_location_features=False))
    for kind, n_img, loc in (("Caption", 1, 0), ("CaptionLoc", 1, 7), ("CaptionGlobalLoc", 2, 1)):
        m, crit, _ = build_product_model(kind)
        assert m.__class__.__name__ == kind and m.n_images == n_img and m.loc_in_dim == loc
        assert isinstance(crit, torch.nn.CrossEntropyLoss)
        frozen = [n for n, p in m.named_parameters() if not p.requires_grad]
        assert all("backbone.body.conv1" in n or "backbone.body.layer1" in n for n in frozen) and frozen


class _StubTok:
    pad_token, cls_token, sep_token = "[PAD]", "[CLS]", "[SEP]"
    _ids = {"[PAD]": 0, "[CLS]": 101, "[SEP]": 102}

    def convert_tokens_to_ids(self, t):
        return self._ids[t]

    def decode(self, ids, skip_special_tokens=True):
        ids = [int(i) for i in ids]
        return " ".join(f"w{i}" for i in ids if not (skip_special_tokens and i in (0, 101, 102)))

    def batch_decode(self, rows, skip_special_tokens=True):
        return [self.decode(r, skip_special_tokens) for r in rows]

    def encode(self, s):
        return [101] + [int(w[1:]) for w in s.split()] + [102]


def _scripted(script, V=200, with_att=False):
    """Foreign callable emitting script[b][step] at the position being predicted."""
    def stub(*args, return_attention=False):
        caption, cap_mask = args[-2], args[-1]
        B, T = caption.shape
        out = torch.zeros(B, T, V)
        filled = int((~cap_mask[0]).sum())
        for b in range(B):
            out[b, filled - 1, script[b][min(filled - 1, len(script[b]) - 1)]] = 1.0
        if return_attention:
            return out, {"step": filled - 1}
        return out
    stub.eval = lambda: stub
    return stub


def test_greedy_with_att_and_greedy_single_stop_rules():
    """greedy_with_att WRITES the EOS and returns ids without BOS (decode.py:153-163); greedy_single stops before
    writing it and decodes the whole buffer (decode.py:40-50)."""
    from retr_b200.eval_utils import decode
    from retr_b200.models.utils import NestedTensor
    tok = _StubTok()
    s = (NestedTensor(torch.zeros(1, 3, 4, 4), torch.zeros(1, 4, 4, dtype=torch.bool)),)
    ids, atts = decode.greedy_with_att(_scripted({0: [7, 9, 102, 5]}), s, tok, 101, 102, 8, device="cpu")
    assert ids.tolist() == [7, 9, 102] and [a["step"] for a in atts] == [0, 1, 2]
    text, atts = decode.greedy_with_att(_scripted({0: [7, 9, 102, 5]}), s, tok, 101, 102, 8, return_raw=False,
                                        device="cpu")
    assert text == "w7 w9" and len(atts) == 3
    # never emits EOS -> runs to the end of the buffer
    ids, atts = decode.greedy_with_att(_scripted({0: [4]}), s, tok, 101, 102, 6, device="cpu")
    assert ids.tolist() == [4] * 5 and len(atts) == 5
    assert decode.greedy_single(_scripted({0: [7, 9, 102, 5]}), s[0], tok, 101, 102, 8) == "w7 w9"


def test_eval_model_shell_collects_hypotheses_and_references():
    from types import SimpleNamespace

    from retr_b200 import engine
    tok = _StubTok()
    t, tm = torch.zeros(2, 3, 4, 4), torch.zeros(2, 4, 4, dtype=torch.bool)
    batches = [(torch.tensor([11, 12]), t, tm, None, None), (torch.tensor([13, 11]), t, tm, None, None)]

    class Loader(list):
        dataset = SimpleNamespace(annot=[(11, 0, "w5 w6"), (11, 0, "w5"), (12, 0, "w8"), (12, 0, "w8 w8"),
                                         (13, 0, "w1"), (13, 0, "w2")],
                                  return_global_context=False, return_location_features=False)
    seen = {}

    def metrics(ref_list, hyp_list):
        seen.update(ref=ref_list, hyp=hyp_list)
        return {"n": len(hyp_list)}

    model = _scripted({0: [7, 9, 102], 1: [8, 102, 4]})
    out, ids_hyps = engine.eval_model(model, Loader(batches), tok, SimpleNamespace(max_position_embeddings=8),
                                      compute_metrics=metrics)
    assert out == {"n": 4}
    assert ids_hyps == [{"ann_id": 11, "expression": "w7 w9"}, {"ann_id": 12, "expression": "w8"},
                        {"ann_id": 13, "expression": "w7 w9"}, {"ann_id": 11, "expression": "w8"}]
    assert seen["hyp"] == ["w7 w9", "w8", "w7 w9", "w8"]
    assert seen["ref"] == [["w5 w6", "w8", "w1", "w5 w6"], ["w5", "w8 w8", "w2", "w5"]]   # transposed (engine.py:180)
    assert engine.eval_model(model, Loader(batches), tok, SimpleNamespace(max_position_embeddings=8))[0] == {}


def test_greedy_stream_two_contexts_call_order(monkeypatch):
    """Host logic of greedy_stream(contexts=2): the NEXT batch is encoded (asynchronously) on the other context before
    the current one is decoded; outputs come back in batch order; a ragged last batch is fine.  Scripted contexts, no
    device."""
    from retr_b200.eval_utils import decode
    from retr_b200.models.utils import NestedTensor
    m, _, _ = build_product_model("CaptionGlobalLoc")
    log = []

    class FakeEngine:
        def __init__(self, name):
            self.name, self.max_batch, self.cur, self.cap = name, 4, None, None

        def set_encode_cta_cap(self, cap):
            self.cap = cap

        def encode_host(self, t_img, t_mask, g_img, g_mask, loc):
            assert t_mask.dtype == torch.uint8 and t_img.is_contiguous()
            self.cur = int(loc[0, 0])
            log.append((self.name, "encode", self.cur))

        def decode_host(self, bos, eos, max_len):
            log.append((self.name, "decode", self.cur))
            return torch.full((1, max_len), self.cur, dtype=torch.int64), max_len - 1

    engs = [FakeEngine("a"), FakeEngine("b")]
    monkeypatch.setattr(m, "pipeline_engines", lambda batch, n=2: engs)

    def unpinned(model, samples):                      # pin_memory needs a CUDA runtime; the order logic does not
        t, g, loc = model._split_samples(samples)
        return [t.tensors, t.mask.to(torch.uint8), g.tensors, g.mask.to(torch.uint8), loc]
    monkeypatch.setattr(decode, "_stage_host", unpinned)

    def batch(tag, b):
        img, mask = torch.zeros(b, 3, 8, 8), torch.zeros(b, 8, 8, dtype=torch.bool)
        return (NestedTensor(img, mask), NestedTensor(img, mask), torch.full((b, 5), float(tag)))
    outs = list(decode.greedy_stream([batch(0, 2), batch(1, 2), batch(2, 1)], m, max_len=6, bos_token=101, eos_token=102,
                                     contexts=2, encode_cta_cap=100))
    assert [int(o[0, 0]) for o in outs] == [0, 1, 2]
    assert log == [("a", "encode", 0), ("b", "encode", 1), ("a", "decode", 0), ("a", "encode", 2), ("b", "decode", 1),
                   ("a", "decode", 2)]
    assert engs[0].cap == 100 and engs[1].cap == 100
    with pytest.raises(RuntimeError):                  # a later batch larger than the contexts were sized for
        list(decode.greedy_stream([batch(0, 2), batch(1, 5)], m, max_len=6, contexts=2))


def test_check_loc_shapes():
    """a [B,7] loc into CaptionGlobalLoc (5 tokens) or a [B,5] loc into CaptionLoc must raise, as the
    reference's loc_proj Linear would, instead of being read as B*n contiguous floats."""
    from retr_b200 import native
    from retr_b200.runtime import check_loc
    d = native.ModelDesc()
    d.loc_in_dim, d.n_loc = 1, 5
    check_loc(d, torch.zeros(3, 5), 3)
    with pytest.raises(RuntimeError):
        check_loc(d, torch.zeros(3, 7), 3)
    with pytest.raises(ValueError):
        check_loc(d, None, 3)
    d.loc_in_dim, d.n_loc = 7, 7
    with pytest.raises(RuntimeError):
        check_loc(d, torch.zeros(3, 5), 3)
    d.loc_in_dim = 0
    check_loc(d, None, 3)
