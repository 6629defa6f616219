"""Condition producer on the GPU (ds_gemm / ds_layernorm / ds_attention / ds_mask_rows / ds_length_regulate through the C ABI)
against the reference's golden outputs and, at the real width (H = 256, 4 layers, 2 heads: configs/acoustic.yaml), against the
CPU oracle -- bf16 operands: rel-L2 <= 1e-2 on the condition; then the whole chain tokens -> condition -> 20-step sampler."""
import pytest
import torch

import diffsinger_b200 as ds
from conftest import load_fs2_golden, rel_l2
from oracle.fastspeech_oracle import fs2_condition

pytestmark = pytest.mark.gpu
dev = "cuda"
COND_TOL = 1e-2


def _to(d):
    return {k: v.to(dev) for k, v in d.items()}


@pytest.mark.parametrize("tag", ["a", "b"])
def test_native_condition_matches_reference_golden(tag):
    d, y, sd, hp = load_fs2_golden(tag)
    ds.hparams.clear()
    ds.hparams.update(hp)
    m = ds.FastSpeech2Acoustic(sd["txt_embed.weight"].shape[0])
    m.load_state_dict(sd, strict=True)
    m = m.eval().to(dev)
    g = _to(d)
    kw = {k: v for k, v in g.items() if k not in ("tokens", "mel2ph", "f0")}
    with torch.no_grad():
        out = m(g["tokens"], g["mel2ph"], g["f0"], **kw)
        out2 = m(g["tokens"], g["mel2ph"], g["f0"], **kw)
    assert out.shape == y.shape and torch.isfinite(out).all() and torch.equal(out, out2)
    err = rel_l2(out.cpu(), y)
    print(f"\n[fs2_{tag}] native condition vs reference rel-L2 {err:.3e}")
    assert err < COND_TOL
    # frames past the end of an utterance (mel2ph == 0) carry only the per-frame embeddings, exactly like the reference
    pad_frames = d["mel2ph"] == 0
    assert rel_l2(out.cpu()[pad_frames], y[pad_frames]) < 1e-5


def test_real_width_condition_and_sampler_chain():
    H, heads, layers, vocab = 256, 2, 4, 61
    hp = dict(hidden_size=H, enc_layers=layers, num_heads=heads, enc_ffn_kernel_size=3, ffn_act="gelu", dropout=0.1, use_pos_embed=True,
              use_rope=True, rel_pos=True, use_lang_id=False, use_spk_id=False)
    ds.hparams.clear()
    ds.hparams.update(hp)
    torch.manual_seed(5)
    m = ds.FastSpeech2Acoustic(vocab).eval()
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    B, Tt, T = 1, 143, 1197  # more than one 128-token tile; a long .ds segment
    g = torch.Generator().manual_seed(6)
    tok = torch.randint(1, vocab, (B, Tt), generator=g)
    mel2ph = torch.sort(torch.randint(1, Tt + 1, (B, T), generator=g), dim=1).values
    f0 = torch.rand(B, T, generator=g) * 300 + 80
    ref = fs2_condition(sd, tok, mel2ph, f0, num_heads=heads)
    m = m.to(dev)
    with torch.no_grad():
        cond = m(tok.to(dev), mel2ph.to(dev), f0.to(dev))
    err = rel_l2(cond.cpu(), ref)
    print(f"\n[fs2 H{H} L{layers} T_txt{Tt} T{T}] native condition vs oracle rel-L2 {err:.3e}")
    assert err < COND_TOL
    # ... and on into the sampler: the condition is exactly what GaussianDiffusion.forward takes
    ds.hparams.update(schedule_type="linear", use_shallow_diffusion=False, diff_accelerator="ddim", diff_speedup=50, K_step=1000,
                      K_step_infer=1000, timesteps=1000)
    from oracle.wavenet_oracle import random_state_dict
    dm = ds.GaussianDiffusion(128, 1, 1000, 1000, "wavenet", dict(num_layers=4, num_channels=256, dilation_cycle_length=4),
                              spec_min=[-12.0], spec_max=[0.0])
    dm.denoise_fn.load_state_dict(random_state_dict(128, 1, 4, 256, H, seed=8))
    dm = dm.eval().to(dev)
    with torch.no_grad():
        mel = dm(cond, infer=True)
    assert mel.shape == (B, T, 128) and torch.isfinite(mel).all()


def test_tokens_to_mel_in_one_cuda_graph():
    """with the condition producer native, the whole `.ds` tensors -> condition -> 20-step sampler call is ONE
    CUDA graph (torch.cuda.CUDAGraph around the public calls; inside a capture run_program launches its kernels into the caller's
    graph).  Replays with new inputs in the static tensors match eager calls on the same inputs and noise."""
    H, heads, layers, vocab, M = 256, 2, 4, 61, 128
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, enc_layers=layers, num_heads=heads, enc_ffn_kernel_size=3, ffn_act="gelu", dropout=0.1,
                      use_pos_embed=True, use_rope=True, rel_pos=True, use_lang_id=False, use_spk_id=False, schedule_type="linear",
                      use_shallow_diffusion=False, diff_accelerator="ddim", diff_speedup=50, K_step=1000, K_step_infer=1000,
                      timesteps=1000)
    from oracle.wavenet_oracle import random_state_dict
    torch.manual_seed(9)
    enc = ds.FastSpeech2Acoustic(vocab).eval().to(dev)
    dm = ds.GaussianDiffusion(M, 1, 1000, 1000, "wavenet", dict(num_layers=6, num_channels=256, dilation_cycle_length=4),
                              spec_min=[-12.0], spec_max=[0.0])
    dm.denoise_fn.load_state_dict(random_state_dict(M, 1, 6, 256, H, seed=8))
    dm = dm.eval().to(dev)
    B, Tt, T = 1, 40, 300

    def inputs(seed):
        g = torch.Generator().manual_seed(seed)
        tok = torch.randint(1, vocab, (B, Tt), generator=g)
        mel2ph = torch.sort(torch.randint(1, Tt + 1, (B, T), generator=g), dim=1).values
        return tok.to(dev), mel2ph.to(dev), (torch.rand(B, T, generator=g) * 300 + 80).to(dev), torch.randn(B, 1, M, T, generator=g).to(dev)

    def call(tok, mel2ph, f0, noise):
        dm._noise_override = noise
        try:
            return dm(enc(tok, mel2ph, f0), infer=True)
        finally:
            dm._noise_override = None

    s_tok, s_m2p, s_f0, s_noise = inputs(1)
    with torch.no_grad():
        eager1 = call(s_tok, s_m2p, s_f0, s_noise).clone()  # warm-up: sessions, packed weights, step tables
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            call(s_tok, s_m2p, s_f0, s_noise)
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            s_out = call(s_tok, s_m2p, s_f0, s_noise)
        graph.replay()
        torch.cuda.synchronize()
        assert rel_l2(s_out.cpu(), eager1.cpu()) < 1e-6
        tok2, m2p2, f02, noise2 = inputs(2)
        eager2 = call(tok2, m2p2, f02, noise2).clone()
        for dst, src in ((s_tok, tok2), (s_m2p, m2p2), (s_f0, f02), (s_noise, noise2)):
            dst.copy_(src)
        graph.replay()
        torch.cuda.synchronize()
    assert torch.isfinite(s_out).all() and rel_l2(s_out.cpu(), eager2.cpu()) < 1e-6


def test_segments_pipeline_tokens_to_mel_in_padded_and_ragged_batches():
    """encode_segments / infer_segments: the per-segment loop of inference/ds_acoustic.py:234-244 (tokens -> condition -> sampler) over
    segments of different lengths as padded encoder batches + ragged sampler batches; every segment equals its batch-1 result."""
    from diffsinger_b200.synthetic import random_state_dict
    H, heads, layers, vocab, M = 256, 2, 2, 61, 128
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, enc_layers=layers, num_heads=heads, enc_ffn_kernel_size=3, ffn_act="gelu", dropout=0.1, use_pos_embed=True,
                      use_rope=True, rel_pos=True, use_lang_id=False, use_spk_id=False, use_key_shift_embed=True, use_speed_embed=True,
                      diff_speedup=200, diff_accelerator="ddim", timesteps=1000, K_step=1000, use_shallow_diffusion=False,
                      schedule_type="linear")
    torch.manual_seed(5)
    enc = ds.FastSpeech2Acoustic(vocab).eval().to(dev)
    m = ds.GaussianDiffusion(M, 1, 1000, 1000, "wavenet", dict(num_layers=3, num_channels=256, dilation_cycle_length=3),
                             spec_min=[-12.0], spec_max=[0.0])
    m.denoise_fn.load_state_dict(random_state_dict(M, 1, 3, 256, H, seed=8))
    m = m.eval().to(dev)
    g = torch.Generator().manual_seed(6)
    batches = []
    for Tt, T in ((31, 300), (9, 77), (60, 515), (27, 257), (44, 512)):
        tok = torch.randint(1, vocab, (1, Tt), generator=g)
        mel2ph = torch.sort(torch.randint(1, Tt + 1, (1, T), generator=g), dim=1).values
        batches.append(_to(dict(tokens=tok, mel2ph=mel2ph, f0=torch.rand(1, T, generator=g) * 300 + 80,
                                key_shift=torch.full((1, 1), 2.0), speed=torch.full((1, 1), 1.0))))
    with torch.no_grad():
        conds_alone = [enc(b["tokens"], b["mel2ph"], b["f0"], key_shift=b["key_shift"], speed=b["speed"]) for b in batches]
        torch.manual_seed(3)
        mels_alone = [m(c, infer=True) for c in conds_alone]
    conds = ds.encode_segments(enc, batches)
    for a, b in zip(conds_alone, conds):
        assert a.shape == b.shape
        assert rel_l2(b.cpu(), a.cpu()) < 1e-6, rel_l2(b.cpu(), a.cpu())
    torch.manual_seed(3)
    mels = ds.infer_segments(enc, m, batches)
    for a, b in zip(mels_alone, mels):
        assert a.shape == b.shape
        assert rel_l2(b.cpu(), a.cpu()) < 1e-5
