"""TEST INFRASTRUCTURE ONLY.  Generates tests/golden/*.npz by running the UNMODIFIED reference.

Run in the build container (needs /root/reference):   python -m oracle.make_golden
Weights come from exoground_b200.utils.synth.fill_state_dict (deterministic, name-keyed), inputs
from synth.tan_batch / synth.exo_batch, so fixtures hold only OUTPUTS of the reference modules
(model/tan_model.py, model/exo_ground_model.py, model/keystep_ground_model.py, model/vi_encoder.py,
train/loss.py, train/loss_egoexo4d.py) plus the recipe parameters needed to rebuild the inputs.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_import                      # noqa: E402
from exoground_b200.utils import synth             # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
SUB = 997  # stride for subsampling big gradient tensors


def _np(t):
    return t.detach().to(torch.float32).cpu().numpy()


def _grads(model, store, tag="grad"):
    for name, p in model.named_parameters():
        if p.grad is None:
            continue
        g = p.grad.detach().reshape(-1)
        if tag == "grad":
            store[f"gnorm/{name}"] = _np(g.norm())
        store[f"{tag}/{name}"] = _np(g if g.numel() <= 4096 else g[::SUB])


def _scalars(d, prefix, store):
    for k, v in d.items():
        if isinstance(v, dict):
            continue
        store[f"{prefix}/{k}"] = np.float32(float(v))


def build_tan(r, E, D, d_in, head):
    with ref_import.quiet():
        m = r.tan_model.TemporalAligner(E, D, use_alignability_head=head)
    if d_in != 4096:
        m.video_pre_proj = torch.nn.Linear(d_in, 512, bias=False)
        m.text_pre_proj = torch.nn.Linear(d_in, 512, bias=False)
    sd = synth.fill_state_dict({k: v.shape for k, v in m.state_dict().items()})
    m.load_state_dict(sd)
    return m


def tan_case(r, name, E, D, B, T, N, d_in, np_seed, losses, sub_logits=None):
    m = build_tan(r, E, D, d_in, 1)
    b = synth.tan_batch(B=B, T=T, N=N, d_video=d_in, d_text=d_in)
    store = {"meta": np.array([E, D, B, T, N, d_in, np_seed], dtype=np.int64)}
    inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
    g = torch.Generator().manual_seed(1234)
    abs_pos = torch.rand(B, N, 2, generator=g)
    for i, (th, al, use_abs) in enumerate(losses):
        m.zero_grad()
        np.random.seed(np_seed)
        out = m(b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], None)
        args = synth.tan_args(loss_threshold=th, use_alignability_head=al)
        with ref_import.quiet():
            L = r.loss.get_loss(inp, b["video"], b["text"], b["video_padding_mask"],
                                b["text_padding_mask"], out, args, abs_pos if use_abs else None)
        _scalars(L, f"loss{i}", store)
        store[f"loss{i}/cfg"] = np.array([th, al, use_abs], dtype=np.float32)
        if i == 0:
            # gradients of the smooth loss (no quantile threshold, no median-split BCE labels): the bf16
            # parity tests compare against these, because the discrete keep-sets / labels of the later
            # configurations can legitimately flip under bf16 rounding
            L["loss"].backward(retain_graph=(len(losses) == 1))
            _grads(m, store, "grad0")
            m.zero_grad()
        if i == len(losses) - 1:
            L["loss"].backward()
            _grads(m, store)
    for k, v in out.items():
        a = _np(v)
        if sub_logits and k.startswith("logits_"):
            a = a[:, :, ::sub_logits[0], :, ::sub_logits[1]]
        elif sub_logits and k == "dual_feature_video":
            a = a[:, :, ::sub_logits[0], ::16]
        store[f"out/{k}"] = np.nan_to_num(a, nan=0.0)     # padded text columns are NaN-prone (trap 2)
    # argmax-over-time of last-stage (joint+dual)/2 logits per own (b,k)  (eval_zeroshot_align.py:205,222,237)
    sim = (out["logits_dual"][:, -1] + out["logits_joint"][:, -1]) / 2            # [B,T,B,N]
    own = torch.stack([sim[bb, :, bb, :] for bb in range(B)], 0)                  # [B,T,N]
    store["argmax_t"] = own.argmax(dim=1).numpy().astype(np.int64)
    top2 = own.topk(2, dim=1).values
    store["top2_gap_min"] = _np((top2[:, 0] - top2[:, 1])[~b["text_padding_mask"]].min())
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **store)
    print(name, {k: float(v) for k, v in store.items() if k.endswith("/loss")})


def agree_case(r, name, E=2, D=3, B=5, T=16, N=8, d_in=32, np_seed=5):
    """Self-labelling branch (train/loss.py:91-232): loss dicts of the four temporal_agreement_type modes on an
    'init' model (the reference's in-place -6e4 fill of the own-clip blocks included), gradients of 'keep', and
    one 'cotrain' call fed with separate EMA logits."""
    m = build_tan(r, E, D, d_in, 1)
    b = synth.tan_batch(B=B, T=T, N=N, d_video=d_in, d_text=d_in, seed=4242, annot_inside_video=True)
    store = {"meta": np.array([E, D, B, T, N, d_in, np_seed], dtype=np.int64)}
    inp = {"start": b["start"], "end": b["end"], "text": b["text_str"]}
    for mode in ("keep", "i", "u", "keep-joint"):
        m.zero_grad()
        np.random.seed(np_seed)
        out = m(b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], None)
        args = synth.tan_args(learn_agreement=1, temporal_agreement_type=mode)
        with ref_import.quiet():
            L = r.loss.get_loss(inp, b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], out,
                                args, None)
        _scalars(L, f"loss_{mode}", store)
        if mode == "keep":
            L["loss"].backward()
            _grads(m, store, "grad_keep")
    # cotrain: targets from EMA logits (here: the same model at another positional offset), loss on the online ones
    m.zero_grad()
    np.random.seed(np_seed)
    out = m(b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], None)
    np.random.seed(np_seed + 1)
    with torch.no_grad():
        ema = m(b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], None)
    out["ema-logits_dual"], out["ema-logits_joint"] = ema["logits_dual"], ema["logits_joint"]
    args = synth.tan_args(model="cotrain", learn_agreement=1, temporal_agreement_type="keep")
    with ref_import.quiet():
        L = r.loss.get_loss(inp, b["video"], b["text"], b["video_padding_mask"], b["text_padding_mask"], out, args,
                            None)
    _scalars(L, "loss_cotrain", store)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **store)
    print(name, {k: float(v) for k, v in store.items() if k.endswith("/loss")})


def align_case(r, name, E=2, D=3, d_in=32, seq_len=16):
    """Sliding-window alignment inference: the UNMODIFIED eval/eval_zeroshot_align.py::test_alignment_htm driven
    by the closure of train/main.py:166-184 on synthetic HTM-Align-shaped videos (dataset class and loader patched
    to serve them; sentence strings are indices into the synthetic embedding table)."""
    import types
    ref_root = ref_import.REF
    if os.path.join(ref_root, "eval") not in sys.path:
        sys.path.insert(0, os.path.join(ref_root, "eval"))
    import eval_zeroshot_align as ez
    model = build_tan(r, E, D, d_in, head=0)
    model.random_pos_start = 0
    model.eval()
    recipe = [dict(vlen=70, K=18, seed=11), dict(vlen=45, K=9, seed=12), dict(vlen=130, K=40, seed=13)]
    items = [synth.htm_align_item(d_video=d_in, d_text=d_in, **kw) for kw in recipe]

    class FakeSet:
        def __init__(self, *a, **k):
            pass

    def fake_loader(ds, **k):
        out = []
        for vi, it in enumerate(items):
            out.append({"video": it["video"], "str": [(f"{vi}:{k_}",) for k_ in range(len(it["start"]))],
                        "aligned": torch.tensor([it["aligned"]]), "vid": [f"v{vi}"],
                        "start": torch.tensor([it["start"]]), "end": torch.tensor([it["end"]])})
        return out

    def get_text_visual_sim(video_embed, text_str, interpolate_from=None, abs_text_pos=None):
        vi = int(text_str[0].split(":")[0])
        idx = [int(s.split(":")[1]) for s in text_str]
        te = items[vi]["text"][idx]
        with torch.no_grad():
            j = model.get_text_visual_sim_joint(video_embed, te[None], interpolate_from)
            d = model.get_text_visual_sim_dual(video_embed, te[None], interpolate_from)
        return {"sim": j.transpose(-1, -2) / 0.07, "dual-sim": d.transpose(-1, -2) / 0.07}

    captured = {}
    real_auc = ez.metrics.roc_auc_score

    def spy_auc(tgt, sim):
        captured["tgt"], captured["score"] = np.asarray(tgt), np.asarray(sim)
        return real_auc(tgt, sim)

    ez.HTM_Align, ez.DataLoaderFast = FakeSet, fake_loader
    ez.metrics = types.SimpleNamespace(roc_auc_score=spy_auc)
    args = types.SimpleNamespace(seq_len=seq_len, use_alignability_head=0)
    with ref_import.quiet():
        m = ez.test_alignment_htm(get_text_visual_sim, torch.device("cpu"), args)
    store = {"recall": np.float64(m["Recall"]), "auc": np.float64(m["AUC"]), "score": captured["score"],
             "tgt": captured["tgt"], "seq_len": np.int64(seq_len), "dims": np.array([E, D, d_in]),
             "recipe": np.array([[kw["vlen"], kw["K"], kw["seed"]] for kw in recipe])}
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **store)
    print(name, "Recall", m["Recall"], "AUC", m["AUC"], "sentences", len(captured["score"]))


def sim_case(r, name):
    m = build_tan(r, 2, 2, 32, 1)
    m.eval()
    g = torch.Generator().manual_seed(77)
    video = torch.randn(2, 24, 32, generator=g)
    text = torch.randn(2, 5, 32, generator=g)
    store = {}
    with torch.no_grad():
        store["joint_interp16"] = _np(m.get_text_visual_sim_joint(video, text, interpolate_from=16))
        store["dual_interp16"] = _np(m.get_text_visual_sim_dual(video, text, interpolate_from=16))
        m.random_pos_start = 0
        store["joint_p0"] = _np(m.get_text_visual_sim_joint(video, text))
        store["dual_p0"] = _np(m.get_text_visual_sim_dual(video, text))
        al = m.get_alignability(video, text)
        store["align_dual"] = _np(al["alignability-dual"])
        store["align_joint"] = _np(al["alignability-joint"])
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **store)
    print(name, "ok")


def exo_case(r, name):
    Dv, Dt, B, T, N = 48, 64, 3, 16, 8
    b = synth.exo_batch(B=B, T=T, N=N, d_video=Dv, d_text=Dt, d_ego=Dv)
    store = {"meta": np.array([2, 2, B, T, N, Dv, Dt, 3], dtype=np.int64)}
    with ref_import.quiet():
        m = r.exo_ground_model.ExoGroundingTransformer(2, 2, video_embed_dim=Dv, text_embed_dim=Dt,
                                                       use_distill_nce_loss=True)
    m.load_state_dict(synth.fill_state_dict({k: v.shape for k, v in m.state_dict().items()}))
    np.random.seed(3)
    out = m(b["video_features"], b["narration_features"], b["video_padding_mask"],
            b["narration_padding_mask"], egocentric_video_embed=b["ego_video_features"][:, 0])
    for k, v in out.items():
        store[f"out/{k}"] = _np(v)
    args = synth.exo_args(use_distill_nce_loss=True, same_view_negative=True)
    L, iou = r.loss_egoexo4d.get_loss(b, out, b["narration_padding_mask"], args)
    _scalars(L, "loss", store)
    store["iou"] = _np(iou)
    L["loss"].backward()
    _grads(m, store)
    # start/end parametrisation + 2-way InfoNCE, forward only
    args2 = synth.exo_args(use_distill_nce_loss=True, use_center_duration=False)
    L2, _ = r.loss_egoexo4d.get_loss(b, {k: v.detach() for k, v in out.items()}, b["narration_padding_mask"], args2)
    _scalars(L2, "loss_se", store)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **store)
    print(name, float(L["loss"]), float(L2["loss"]))

    # GroundingModel with a frozen ViewInvariantMLP front-end
    store = {}
    with ref_import.quiet():
        vi = r.vi_encoder.ViewInvariantMLP(video_embed_dim=Dv)
        vi.load_state_dict(synth.fill_state_dict({k: v.shape for k, v in vi.state_dict().items()}, seed=7))
        gm = r.keystep_ground_model.GroundingModel(2, 2, video_embed_dim=Dv, text_embed_dim=Dt,
                                                   use_distill_nce_loss=True, vi_encoder=vi)
    gsd = synth.fill_state_dict({k: v.shape for k, v in gm.state_dict().items()
                                 if not k.startswith("vi_encoder.")})
    gm.load_state_dict({**gsd, **{"vi_encoder." + k: v for k, v in vi.state_dict().items()}})
    np.random.seed(3)
    out = gm(b["video_features"], b["narration_features"], b["video_padding_mask"],
             b["narration_padding_mask"], egocentric_video_embed=b["ego_video_features"][:, 0])
    for k, v in out.items():
        store[f"out/{k}"] = _np(v)
    store["vi/high_dim_features"] = _np(vi(b["video_features"], None, None, None)["high_dim_features"])
    (out["interval_preds"].sum() + out["distill_infonce_loss"]).backward()
    _grads(gm, store)
    np.savez_compressed(os.path.join(OUT, "keystep_small.npz"), **store)
    print("keystep_small ok")


def audio_pairwise_case(r, name):
    """Audio branch without decoder (exo_ground_model.py:135-139,347-350; random positional start on, so the
    draw order of np.random is pinned too) + the pairwise multi-view InfoNCE of GroundingModel
    (keystep_ground_model.py:177-179,251-307), whose value and gradient are identically zero."""
    Dv, Dt, Da, B, T, N, V = 48, 64, 40, 3, 16, 8, 2
    b = synth.exo_batch(B=B, T=T, N=N, d_video=Dv, d_text=Dt)
    g = torch.Generator().manual_seed(11)
    audio = torch.randn(B, T, Da, generator=g)
    amask = torch.zeros(B, T, dtype=torch.bool)
    amask[1, T - 5:] = True
    store = {"meta": np.array([2, 1, B, T, N, Dv, Dt, Da, V, 4], dtype=np.int64), "audio": _np(audio),
             "amask": amask.numpy()}
    with ref_import.quiet():
        m = r.exo_ground_model.ExoGroundingTransformer(2, 1, use_decoder=False, use_audio=True, video_embed_dim=Dv,
                                                       text_embed_dim=Dt, audio_embed_dim=Da, use_text_pos_enc=1)
    m.load_state_dict(synth.fill_state_dict({k: v.shape for k, v in m.state_dict().items()}))
    np.random.seed(4)
    out = m(b["video_features"], b["narration_features"], b["video_padding_mask"], b["narration_padding_mask"],
            audio_embed=audio, audio_padding_mask=amask)
    for k, v in out.items():
        store[f"out/{k}"] = _np(v)
    w = torch.linspace(-1, 1, out["interval_preds"].numel()).view_as(out["interval_preds"])
    ((out["interval_preds"] * w).sum() + out["low_dim_features"].square().mean()).backward()
    _grads(m, store)
    # pairwise InfoNCE through GroundingModel, both mask modes
    view_mask = torch.ones(B, T, dtype=torch.bool)
    view_mask[0, T // 2 + 3:] = False
    store["view_mask"] = view_mask.numpy()
    for mode in ("all", "unmasked"):
        with ref_import.quiet():
            gm = r.keystep_ground_model.GroundingModel(2, 1, video_embed_dim=Dv, text_embed_dim=Dt, multi_view=True,
                                                       num_max_views=V, use_pairwise_distill_nce_loss=True,
                                                       pairwise_distill_mode=mode, random_pos_start=0)
        gm.load_state_dict(synth.fill_state_dict({k: v.shape for k, v in gm.state_dict().items()}))
        out = gm(b["video_features"], b["narration_features"], b["video_padding_mask"],
                 b["narration_padding_mask"], view_mask=view_mask)
        store[f"pair_{mode}/interval_preds"] = _np(out["interval_preds"])
        store[f"pair_{mode}/loss"] = _np(out["distill_infonce_loss"])
        store[f"pair_{mode}/requires_grad"] = np.array(out["distill_infonce_loss"].requires_grad)
        (out["interval_preds"].sum() + out["distill_infonce_loss"]).backward()
        store[f"pair_{mode}/proj_grad_absmax"] = _np(gm.exo_feature_proj.weight.grad.abs().max())
        store[f"pair_{mode}/head_grad"] = _np(gm.grounding_head.weight.grad)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **store)
    print(name, "ok", {k: float(v) for k, v in store.items() if k.endswith("/loss")})


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    r = ref_import.load()
    # known answer in the reference (train/loss.py:19-20)
    c = r.loss.circulant(torch.tensor([0, 1, 2]), 0)
    assert c.tolist() == [[0, 1, 2], [2, 0, 1], [1, 2, 0]]
    tan_case(r, "tan_small", 2, 3, 3, 16, 8, 32, 5,
             [(0.0, 0, 0), (0.3, 0, 0), (0.0, 1, 1), (0.5, 1, 0)])
    sim_case(r, "tan_sim")
    agree_case(r, "tan_agree")
    align_case(r, "tan_align")
    exo_case(r, "exo_small")
    audio_pairwise_case(r, "exo_audio")
    tan_case(r, "tan_cfg1", 6, 6, 8, 64, 32, 1024, 888, [(0.0, 0, 0), (0.5, 1, 0)], sub_logits=(8, 4))


if __name__ == "__main__":
    main()
