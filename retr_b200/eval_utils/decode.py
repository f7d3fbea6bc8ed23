"""Greedy decoding with the reference's signatures (eval_utils/decode.py).

``greedy`` keeps the reference contract — ``samples`` tuple + an opaque ``model`` callable in,
``[B, max_len]`` int64 ids on ``device`` out, identical early-return semantics — but when the
model is a retr_b200 caption model the whole loop runs device-resident: encode once, then
``max_len-1`` KV-cached single-token steps replayed from one captured CUDA graph, instead of
re-running the full model (both ResNet passes included) per token (reference
eval_utils/decode.py:68-79).  A foreign callable gets the literal per-token loop.
"""
import torch

from ..models.caption import _CaptionBase


def create_caption_and_mask(start_token, max_length, batch_size=1):
    """BOS-initialised id buffer and its padding mask (True = not yet written).
    reference eval_utils/decode.py:20-27."""
    caption = torch.zeros((batch_size, max_length), dtype=torch.long)
    mask = torch.ones((batch_size, max_length), dtype=torch.bool)
    caption[:, 0] = start_token
    mask[:, 0] = False
    return caption, mask


def _literal_greedy(samples, model, max_len, device, bos_token, eos_token):
    caption, cap_mask = create_caption_and_mask(bos_token, max_len, samples[0].shape[0])
    samples = [s.to(device) for s in samples]
    caption, cap_mask = caption.to(device), cap_mask.to(device)
    finished = torch.zeros(caption.shape[0], dtype=torch.bool, device=device)
    for i in range(max_len - 1):
        ids = model(*samples, caption, cap_mask)[:, i, :].argmax(dim=-1)
        finished |= ids == eos_token
        if bool(finished.all()):
            break
        caption[:, i + 1] = ids
        cap_mask[:, i + 1] = False
    return caption


def greedy(samples, model, max_len=20, device="auto", bos_token=1, eos_token=2):
    """Batch greedy decoding — reference eval_utils/decode.py:53-81."""
    if device == "auto":
        device = "cuda" if torch.cuda.is_available() else "cpu"
    if isinstance(model, _CaptionBase):
        if torch.device(device).type != "cuda":
            raise RuntimeError("retr_b200 models decode on a CUDA device only (no CPU path)")
        samples = [s.to(device) for s in samples]
        ids, _steps = model.greedy_native(samples, max_len, bos_token, eos_token)
        return ids
    return _literal_greedy(samples, model, max_len, device, bos_token, eos_token)


def _stage_host(model, samples):
    t, g, loc = model._split_samples(samples)
    host = [t.tensors, t.mask.to(torch.uint8), None if g is None else g.tensors,
            None if g is None else g.mask.to(torch.uint8), loc]
    out = []
    for x in host:
        if x is not None:
            x = x.float() if x.dtype not in (torch.uint8, torch.float32) else x
            x = x.contiguous()
            x = x if x.is_pinned() else x.pin_memory()
        out.append(x)
    return out


def greedy_stream(batches, model, max_len=20, bos_token=1, eos_token=2, contexts=1, encode_cta_cap=None):
    """``greedy`` over an ITERABLE of host sample tuples (what a DataLoader yields after ``pack_encoder_inputs``);
    yields one ``[B, max_len]`` int64 HOST tensor per batch, equal to ``greedy(batch, model, ...).cpu()``.  The
    reference's loop handles one batch at a time (``samples.to(device)`` at eval_utils/decode.py:62 inside
    engine.py:138-146); here batches overlap:

    * ``contexts=1`` (default): one context; the H2D copies of batch i+1 overlap batch i (``retr_prefetch_host``).
    * ``contexts=2``: two native contexts, each on its own stream.  While batch i runs its 127 decode steps on one
      context, the images of batch i+1 cross PCIe and go through the backbone and encoder on the other
      (``retr_encode_host`` / ``retr_decode_host``).  ``encode_cta_cap`` bounds the encode-phase GEMM grids so decode
      kernels find free SMs.  Measured at B=256: 72.4 ms per batch against 73.4 ms for ``contexts=1`` — the decode
      step's 8-CTA cluster kernels cannot be placed while persistent GEMM CTAs sit in every GPC, so only the encoder
      phase really overlaps; kept as an option, not the default."""
    if not isinstance(model, _CaptionBase):
        raise RuntimeError("greedy_stream needs a retr_b200 caption model (use greedy for a foreign callable)")
    it = iter(batches)
    try:
        cur = _stage_host(model, next(it))
    except StopIteration:
        return
    if contexts >= 2:
        engs = model.pipeline_engines(cur[0].shape[0], 2)
        cap = encode_cta_cap if encode_cta_cap is not None else getattr(model, "encode_cta_cap", None)
        if cap is not None:
            for e in engs:
                e.set_encode_cta_cap(cap)
        i = 0
        engs[0].encode_host(*cur)
        while cur is not None:
            try:
                nxt = _stage_host(model, next(it))
            except StopIteration:
                nxt = None
            if nxt is not None:
                if nxt[0].shape[0] > engs[0].max_batch:
                    raise RuntimeError("greedy_stream: a later batch is larger than the first; set model.max_batch")
                engs[(i + 1) % 2].encode_host(*nxt)       # asynchronous: runs under the decode loop below
            ids, _ = engs[i % 2].decode_host(bos_token, eos_token, max_len)
            yield ids
            cur = nxt
            i += 1
        return
    eng = model.engine(cur[0].shape[0])
    eng.prefetch_host(*cur)
    while cur is not None:
        try:
            nxt = _stage_host(model, next(it))
        except StopIteration:
            nxt = None
        if nxt is not None:
            if nxt[0].shape[0] > eng.max_batch:
                raise RuntimeError("greedy_stream: a later batch is larger than the first; set model.max_batch")
            eng.prefetch_host(*nxt)          # second staging set; crosses PCIe while `cur` is decoded
        ids, _ = eng.generate_host(*cur, bos_token, eos_token, max_len)
        yield ids
        cur = nxt


def prune_cap_ids(idx_seqs, clean=True, pad_token=0, bos_token=1, eos_token=2):
    """Cut each row after its first EOS; ``clean`` drops every pad/bos/eos id.
    reference eval_utils/decode.py:84-101."""
    special = (pad_token, bos_token, eos_token)
    out = []
    for seq in idx_seqs:
        seq = list(seq)
        if eos_token in seq:
            seq = seq[:seq.index(eos_token) + 1]
        out.append([t for t in seq if t not in special] if clean else seq)
    return out


def idx2sents(idx_seqs, tokenizer, skip_special_tokens=True):
    """reference eval_utils/decode.py:104-109."""
    return tokenizer.batch_decode(idx_seqs, skip_special_tokens=skip_special_tokens)


def greedy_decoding(samples, model, tokenizer, max_len=20, clean=True, pad_token=0, bos_token=1,
                    eos_token=2, device="auto"):
    """ids -> pruned ids -> strings; reference eval_utils/decode.py:112-128."""
    ids = greedy(samples, model, max_len=max_len, bos_token=bos_token, eos_token=eos_token, device=device)
    rows = ids.cpu().numpy().tolist()
    return idx2sents(prune_cap_ids(rows, clean=clean, pad_token=pad_token, bos_token=bos_token,
                                   eos_token=eos_token), tokenizer)


def greedy_with_att(model, sample, tokenizer, start_token=1, end_token=2, max_pos_embeddings=128, return_raw=True,
                    device="auto"):
    """Single-sample greedy decoding that also returns, per generated token, the attention dict of the
    forward that produced it — reference eval_utils/decode.py:131-167 (literal: one full
    ``return_attention=True`` forward per token; EOS IS written before the loop stops)."""
    if device == "auto":
        device = "cuda" if torch.cuda.is_available() else "cpu"
    caption, cap_mask = create_caption_and_mask(start_token, max_pos_embeddings)
    sample = [s.to(device) for s in sample]
    caption, cap_mask = caption.to(device), cap_mask.to(device)
    atts = []
    with torch.no_grad():
        model.eval()
        for i in range(max_pos_embeddings - 1):
            predictions, att = model(*sample, caption, cap_mask, return_attention=True)
            predicted_id = torch.argmax(predictions[:, i, :], axis=-1)
            caption[:, i + 1] = predicted_id[0]
            cap_mask[:, i + 1] = False
            atts.append(att)
            if predicted_id[0] == end_token:
                break
    token_ids = caption[0][~cap_mask[0]][1:]            # exclude BOS
    if return_raw:
        return token_ids, atts
    return tokenizer.decode(token_ids, skip_special_tokens=True), atts


def greedy_single(model, image, tokenizer, start_token, end_token, max_pos_embeddings):
    """Greedy decoding of one image with the target-only ``Caption`` model — reference
    eval_utils/decode.py:30-50: stops WITHOUT writing EOS and decodes the whole id buffer.  A retr_b200 model runs the
    device-resident KV-cached loop (for a batch of one the early-return rule of ``greedy`` is this very rule);
    a foreign callable gets the literal loop."""
    if isinstance(model, _CaptionBase):
        device = next(model.parameters()).device
        with torch.no_grad():
            model.eval()
            ids, _ = model.greedy_native([image.to(device) if hasattr(image, "to") else image], max_pos_embeddings,
                                         start_token, end_token)
        return tokenizer.decode(ids[0].cpu(), skip_special_tokens=True)
    caption, cap_mask = create_caption_and_mask(start_token, max_pos_embeddings)
    with torch.no_grad():
        model.eval()
        for i in range(max_pos_embeddings - 1):
            predicted_id = torch.argmax(model(image, caption, cap_mask)[:, i, :], axis=-1)
            if predicted_id[0] == end_token:
                break
            caption[:, i + 1] = predicted_id[0]
            cap_mask[:, i + 1] = False
    return tokenizer.decode(caption[0], skip_special_tokens=True)
