"""Generate golden vectors by running the UNMODIFIED reference in the build container.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

Imports ``msstftd``, ``pkl_file_model_updated`` and ``pkl_file_models`` from ``/root/reference``
(read-only; not available on the GPU box, which is why the outputs are committed as ``*.npz``),
feeds them seeded inputs and stores inputs + outputs:

* ``stft_disc.npz``   -- ``DiscriminatorSTFT.spec_transform`` + the ``cat(real, imag)`` / rearrange
  of ``msstftd.py:151-153`` for the three default scales (``msstftd.py:174-175``), plus the logit
  shapes of ``MultiScaleSTFTDiscriminator(filters=32)`` on ``(2, 1, 24000)``
  (known answer of ``testing_scripts.ipynb:2021-2023``).
* ``dpca.npz``        -- ``data_pca`` (``pkl_file_model_updated.py:37-42``), ``apply_pca``
  (``pkl_file_models.py:224-228``) and the DPCA blocks of ``ResE2D1.forward`` (``:486-511``) and
  ``ResE4D1.forward`` (``:394-427``) run through the reference's own ``forward`` with the encoders /
  decoder replaced by pass-through stubs (so the latents are the seeded inputs and the tensor handed
  to the decoder is captured).
* ``spec_losses.npz`` -- the loss dictionary / spectral SNR of ``ResE2D1.forward`` (``:520-540``).
* ``sgmse_spec.npz``  -- the SGMSE+ front end ``pad_spec(spec_fwd(stft(sig)))`` (``sgmse/sgmse/data_module.py:162-175,212-214``,
  ``sgmse/sgmse/util/other.py:76-90``), the three functions compiled from the reference source with ``ast``.
* ``bottleneck.npz``  -- the split / alpha-mix / truncate bottleneck of ``SpectralResE2D2`` / ``SpectralResE4D1``
  (``pkl_file_models.py:1108-1127``, ``:692-718``) through their own ``forward``, and the commented-out DPCA blocks of
  ``:302-332`` / ``:636-679`` executed as written.
* ``task_aware_tail.npz`` -- ``train_pkl_file_task_aware.py:312-364`` executed as written on seeded inputs with a stub discriminator.
* ``var_topk.npz``    -- variance top-k of ``pkl_file_models.py:901-907`` through
  ``SpectralResE1D1.forward`` with stubbed encoder / decoder.
"""
import os
import sys
import warnings

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.dont_write_bytecode = True
sys.path.insert(0, REF)
warnings.filterwarnings("ignore")

import msstftd  # noqa: E402
import pkl_file_model_updated as pmu  # noqa: E402
import pkl_file_models as pm  # noqa: E402


class _Enc(torch.nn.Module):
    """Pass-through encoder stub: the 'observation' already is the latent, shaped [B,1,1,d]."""
    def forward(self, obs):
        return obs.reshape(obs.shape[0], -1), None


class _Dec(torch.nn.Module):
    def __init__(self, shape_fn):
        super().__init__()
        self.shape_fn = shape_fn
        self.seen = None

    def forward(self, z):
        self.seen = z.detach().clone()
        return self.shape_fn(z)


def _bare(cls):
    m = cls.__new__(cls)
    torch.nn.Module.__init__(m)
    return m


def gen_stft():
    torch.manual_seed(20240)
    B, T = 2, 6000
    x = 0.3 * torch.randn(B, 1, T)
    x[:, 0] += 0.5 * torch.sin(2 * np.pi * 440.0 * torch.arange(T) / 16000.0)
    out = {"x": x.numpy()}
    disc = msstftd.MultiScaleSTFTDiscriminator(filters=4)
    for i, d in enumerate(disc.discriminators):
        z = d.spec_transform(x)                                  # msstftd.py:151
        out[f"complex_{i}"] = z.numpy()
        zz = torch.cat([z.real, z.imag], dim=1)                  # msstftd.py:152
        out[f"planes_{i}"] = zz.permute(0, 1, 3, 2).contiguous().numpy()   # 'b c w t -> b c t w'
        out[f"cfg_{i}"] = np.array([d.n_fft, d.hop_length, d.win_length])
    torch.manual_seed(1)
    d32 = msstftd.MultiScaleSTFTDiscriminator(filters=32)
    logits, fmaps = d32(torch.randn(2, 1, 24000))
    out["kat_logit_shapes"] = np.array([list(l.shape) for l in logits])
    out["kat_n_fmaps"] = np.array([len(f) for f in fmaps])
    np.savez_compressed(os.path.join(HERE, "stft_disc.npz"), **out)


def _latents(B, d, seed, decay=0.93):
    g = torch.Generator().manual_seed(seed)
    sig = decay ** torch.arange(d, dtype=torch.float32)
    Q, _ = torch.linalg.qr(torch.randn(d, d, generator=g))
    return ((torch.randn(B, d, generator=g) * sig) @ Q.T) * (1.0 + 0.01 * (seed % 7)) + 0.3


def gen_dpca():
    out = {}
    # data_pca / apply_pca, tall and wide (B < d exercises the rank-deficient thin SVD)
    for tag, (B, d) in {"tall": (96, 32), "wide": (12, 32)}.items():
        z = _latents(B, d, 11 if tag == "tall" else 12)
        s, v, mu = pmu.data_pca(z)
        s2, v2, mu2 = pm.data_pca(z)
        assert torch.equal(s, s2) and torch.equal(mu, mu2)
        out[f"pca_{tag}_z"] = z.numpy(); out[f"pca_{tag}_s"] = s.numpy()
        out[f"pca_{tag}_v"] = v.numpy(); out[f"pca_{tag}_mu"] = mu.numpy()
        out[f"pca_{tag}_apply8"] = pm.apply_pca(z, 8).numpy()

    # two-view DPCA block through ResE2D1.forward
    B, d = 64, 32
    z1, z2 = _latents(B, d, 21), _latents(B, d, 22)
    m = _bare(pmu.ResE2D1)
    m.enc1, m.enc2 = _Enc(), _Enc()
    m.dec = _Dec(lambda z: torch.zeros(z.shape[0], 2, 1, d))
    torch.manual_seed(77)
    state = torch.random.get_rng_state()
    dim_p = torch.randint(8, int(2 * d / 2), (1,)).item()        # same draw as :489
    torch.random.set_rng_state(state)
    m.forward(z1.reshape(B, 1, 1, d), z2.reshape(B, 1, 1, d), True)
    out["e2_z1"], out["e2_z2"] = z1.numpy(), z2.numpy()
    out["e2_dim_p"] = np.array(dim_p)
    out["e2_out"] = m.dec.seen.numpy()                            # cat(z1', z2') handed to the decoder

    # four-view DPCA block through ResE4D1.forward
    zs = [_latents(B, d, 31 + i) for i in range(4)]
    m4 = _bare(pmu.ResE4D1)
    m4.enc1, m4.enc2, m4.enc3, m4.enc4 = _Enc(), _Enc(), _Enc(), _Enc()
    m4.dec = _Dec(lambda z: torch.zeros(z.shape[0], 2, 2, d))
    torch.manual_seed(78)
    state = torch.random.get_rng_state()
    dim_p4 = torch.randint(8, int(4 * d / 2), (1,)).item()       # :398
    torch.random.set_rng_state(state)
    m4.forward(*[z.reshape(B, 2, 1, d // 2) for z in zs], True)
    for i, z in enumerate(zs):
        out[f"e4_z{i + 1}"] = z.numpy()
    out["e4_dim_p"] = np.array(dim_p4)
    out["e4_out"] = m4.dec.seen.numpy()
    np.savez_compressed(os.path.join(HERE, "dpca.npz"), **out)


def gen_spec_losses():
    torch.manual_seed(5)
    B, K, F = 3, 65, 20
    obs1, obs2 = torch.rand(B, 1, K, F) * 2.0, (torch.rand(B, 1, K, F) - 0.5) * 6.0
    dec_out = torch.cat((obs1, obs2), 1) + 0.1 * torch.randn(B, 2, K, F)
    m = _bare(pmu.ResE2D1)

    class _E(torch.nn.Module):
        def forward(self, o):
            return o.reshape(o.shape[0], -1)[:, :16], None
    m.enc1, m.enc2 = _E(), _E()
    m.dec = _Dec(lambda z: dec_out)
    r = m.forward(obs1, obs2, False)
    obs_dec, mse_mean, nuc, _, cos, sd, snr, psnr = r
    obs = torch.cat((obs1, obs2), 1)
    np.savez_compressed(os.path.join(HERE, "spec_losses.npz"), obs=obs.numpy(), dec=obs_dec.numpy(),
                        mse_mean=mse_mean.numpy(), magnitude_loss=sd["magnitude_loss"].numpy(),
                        phase_loss=sd["phase_loss"].numpy(), total_loss=sd["total_loss"].numpy(),
                        spec_snr=snr.numpy(), nuc=nuc.numpy(), z_sample=m.dec.seen.numpy())


def gen_var_topk():
    B, d, k = 48, 64, 8
    z = _latents(B, d, 41, decay=0.9)
    m = _bare(pm.SpectralResE1D1)
    m.total_features_after = k
    m.enc = _Enc()
    m.dec = _Dec(lambda zz: torch.zeros(zz.shape[0], 2, 1, d // 2))
    obs = {"magnitude": z[:, : d // 2].reshape(B, 1, d // 2), "phase": z[:, d // 2:].reshape(B, 1, d // 2)}
    m.forward(obs, None, True)
    # the encoder stub flattens stack([mag, phase], 1) -> [B, d] == z
    np.savez_compressed(os.path.join(HERE, "var_topk.npz"), z=z.numpy(), k=np.array(k),
                        z_sel=m.dec.seen.numpy())


def _reference_function(path, name, patches=()):
    """Compile ONE function of a reference script that cannot be imported as a module (``train_pkl_file_task_aware.py``
    imports librosa / speechbrain and downloads checkpoints at import time): its source is read from the reference
    checkout at generation time, never stored in this repository.  ``patches`` are (old, new) string replacements applied
    to that source -- only used to turn the hard-coded ``"cuda:0"`` of ``:189`` into ``"cpu"`` (no GPU in this container)."""
    import ast
    src = open(os.path.join(REF, path)).read()
    node = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == name)
    text = ast.get_source_segment(src, node)
    for old, new in patches:
        assert old in text
        text = text.replace(old, new)
    ns = {"torch": torch, "np": np}
    exec(compile(text, f"{path}:{name}", "exec"), ns)
    return ns[name]


def gen_istft():
    """``batch_reconstruct_waveform`` / ``reconstruct_waveform`` of ``train_pkl_file_task_aware.py:174-226`` on a seeded
    magnitude / phase batch shaped like the decoder output (K = 1025)."""
    brw = _reference_function("train_pkl_file_task_aware.py", "batch_reconstruct_waveform", [('"cuda:0"', '"cpu"')])
    rw = _reference_function("train_pkl_file_task_aware.py", "reconstruct_waveform", [('"cuda:0"', '"cpu"')])
    g = torch.Generator().manual_seed(909)
    B, K, F = 3, 1025, 14
    mag = torch.rand(B, K, F, generator=g) * torch.linspace(2.0, 0.05, K)[None, :, None]
    phase = (torch.rand(B, K, F, generator=g) - 0.5) * 2.0 * np.pi * 1.2          # a little beyond [-pi, pi] like a raw decoder
    wav = brw(mag, phase, 16000)
    one = rw(mag[0], phase[0], 16000)
    np.savez_compressed(os.path.join(HERE, "istft.npz"), mag=mag.numpy(), phase=phase.numpy(), wav=wav.numpy(),
                        one=one.numpy())


def gen_cosine():
    """``cos_loss`` of ``ResE4D1.forward`` (``pkl_file_model_updated.py:429-430``) and ``ResE2D1.forward`` (``:513-514``),
    evaluation branch (no DPCA), encoders / decoder stubbed as in ``gen_dpca``."""
    B, d = 40, 24
    out = {}
    zs = [_latents(B, d, 51 + i) - 0.3 for i in range(4)]
    m4 = _bare(pmu.ResE4D1)
    m4.enc1, m4.enc2, m4.enc3, m4.enc4 = _Enc(), _Enc(), _Enc(), _Enc()
    m4.dec = _Dec(lambda z: torch.zeros(z.shape[0], 2, 2, d))
    r4 = m4.forward(*[z.reshape(B, 2, 1, d // 2) for z in zs], False)
    for i, z in enumerate(zs):
        out[f"c4_z{i + 1}"] = z.numpy()
    out["c4_cos"] = r4[4].detach().numpy()
    m2 = _bare(pmu.ResE2D1)
    m2.enc1, m2.enc2 = _Enc(), _Enc()
    m2.dec = _Dec(lambda z: torch.zeros(z.shape[0], 2, 1, d))
    r2 = m2.forward(zs[0].reshape(B, 1, 1, d), zs[1].reshape(B, 1, 1, d), False)
    out["c2_cos"] = r2[4].detach().numpy()
    np.savez_compressed(os.path.join(HERE, "cosine.npz"), **out)


def _reference_method(path, cls, name):
    """Like ``_reference_function`` for ONE method of a class in a module that cannot be imported (``sgmse.data_module``
    needs pytorch_lightning): the method's source is read from the reference at generation time and compiled as a plain
    function taking ``self``."""
    import ast
    import textwrap
    src = open(os.path.join(REF, path)).read()
    cnode = next(n for n in ast.parse(src).body if isinstance(n, ast.ClassDef) and n.name == cls)
    node = next(n for n in cnode.body if isinstance(n, ast.FunctionDef) and n.name == name)
    text = textwrap.dedent(ast.get_source_segment(src, node, padded=True))
    ns = {"torch": torch, "np": np}
    exec(compile(text, f"{path}:{cls}.{name}", "exec"), ns)
    return ns[name]


def gen_sgmse():
    """The SGMSE+ front end: ``SpecsDataModule.stft`` / ``spec_fwd`` (``sgmse/sgmse/data_module.py:162-175,212-214``) with the
    module's default hyper-parameters (``:109-118,123-125``), the window of ``get_window`` (``:13-17``), then ``pad_spec``
    (``sgmse/sgmse/util/other.py:76-90``) on the ``[B, 1, K, F]`` tensor as ``sgmse/test_single.py:148-149`` does."""
    import types
    dm = "sgmse/sgmse/data_module.py"
    get_window = _reference_function(dm, "get_window")
    self = types.SimpleNamespace(n_fft=510, hop_length=128, window=get_window("hann", 510), windows={}, spec_factor=0.15,
                                 spec_abs_exponent=0.5, transform_type="exponent")
    for name in ("spec_fwd", "_get_window", "stft"):
        setattr(self, name, types.MethodType(_reference_method(dm, "SpecsDataModule", name), self))
    self.stft_kwargs = {"n_fft": 510, "hop_length": 128, "window": self.window, "center": True, "return_complex": True}   # :192-199
    pad_spec = _reference_function("sgmse/sgmse/util/other.py", "pad_spec")
    g = torch.Generator().manual_seed(510)
    sig = torch.randn(2, 6000, generator=g) * 0.1 + 0.05 * torch.sin(torch.arange(6000) * 0.05)[None]
    Y = self.spec_fwd(self.stft(sig))                                  # [2, 256, 47]
    Yp = pad_spec(Y.unsqueeze(1))                                      # [2, 1, 256, 64]
    np.savez_compressed(os.path.join(HERE, "sgmse_spec.npz"), sig=sig.numpy(), F=np.array(Y.shape[-1]), Yp=Yp.squeeze(1).numpy())


def _reference_lines(path, first, last, first_has, last_has, uncomment=False):
    """Source lines [first, last] (1-based, inclusive) of a reference file, dedented -- for blocks that sit in the middle of
    a function that cannot be imported or called (``task_aware`` needs wav2vec2 + a processor) or that upstream has
    commented out (the DPCA blocks of ``pkl_file_models.py``; ``uncomment`` strips ONE leading ``# `` per line).  The two
    ``*_has`` strings anchor the range so that a moved block fails loudly.  Nothing is stored in this repository."""
    import textwrap
    lines = open(os.path.join(REF, path)).read().split("\n")[first - 1:last]
    assert first_has in lines[0] and last_has in lines[-1], (lines[0], lines[-1])
    if uncomment:
        out = []
        for ln in lines:
            t = ln.lstrip()
            assert t == "" or t.startswith("#"), ln
            out.append(t[2:] if t.startswith("# ") else t[1:])
        lines = out
    return textwrap.dedent("\n".join(lines))


def gen_bottleneck():
    """Row a9.  (1) The bottleneck that is live in ``pkl_file_models.py``: ``SpectralResE2D2.forward`` (``:1108-1127``) and
    ``SpectralResE4D1.forward`` (``:692-718``) with pass-through encoder stubs, the tensors handed to the decoders captured;
    alpha is re-drawn from the same seed.  (2) The DPCA blocks upstream has commented out at the same place, executed as
    written: E2D1 ``:302-332`` (``dim_p = num_features / 2``) and E4D1 ``:636-679`` (``dim_p = randint(8, num_features/2)``)."""
    import contextlib
    import io
    import types
    out = {}
    B, d, tfa = 48, 32, 24
    z1, z2 = _latents(B, d, 61), _latents(B, d, 62)

    def obs_of(z):          # stack([magnitude, phase], 1).float() -> the stub encoder flattens it back to z
        return {"magnitude": z[:, :d // 2].reshape(B, 1, d // 2), "phase": z[:, d // 2:].reshape(B, 1, d // 2)}

    class _EncP(torch.nn.Module):
        def forward(self, obs):
            return obs.reshape(obs.shape[0], -1), None

    m = _bare(pm.SpectralResE2D2)
    m.total_features_after = tfa
    m.enc1, m.enc2 = _EncP(), _EncP()
    m.dec1 = _Dec(lambda z: torch.zeros(B, 2, 1, d // 2))
    m.dec2 = _Dec(lambda z: torch.zeros(B, 2, 1, d // 2))
    m.dimension_info = {}
    for rb in (True, False):
        torch.manual_seed(91)
        alpha = torch.rand(B, 1)
        torch.manual_seed(91)
        r = m.forward(obs_of(z1), obs_of(z2), None, rb)
        tag = "mix" if rb else "nomix"
        out[f"e2d2_{tag}_s1"], out[f"e2d2_{tag}_s2"] = m.dec1.seen.numpy(), m.dec2.seen.numpy()
        if rb:
            out["e2d2_alpha"] = alpha.numpy()
            out["e2d2_after"] = np.array([r[-1]["after_z1"], r[-1]["after_z2"]])
    out["e2d2_z1"], out["e2d2_z2"], out["e2d2_tfa"] = z1.numpy(), z2.numpy(), np.array(tfa)

    zs = [_latents(B, d, 71 + i) for i in range(4)]
    m4 = _bare(pm.SpectralResE4D1)
    m4.total_features_after = tfa
    m4.enc1, m4.enc2, m4.enc3, m4.enc4 = _EncP(), _EncP(), _EncP(), _EncP()
    m4.dec = _Dec(lambda z: torch.zeros(B, 2, 2, d))
    m4.dimension_info = {}
    torch.manual_seed(92)
    m4.forward(*[obs_of(z) for z in zs], None, True)
    out["e4d1_mix"] = m4.dec.seen.numpy()
    for i, z in enumerate(zs):
        out[f"e4d1_z{i + 1}"] = z.numpy()

    # (2) the commented-out DPCA blocks, run as written
    ns = {"torch": torch, "data_pca": pm.data_pca, "self": types.SimpleNamespace(), "random_bottle_neck": True,
          "z1": z1.clone(), "z2": z2.clone(), "num_features": 2 * d}
    code = _reference_lines("pkl_file_models.py", 302, 332, "# if random_bottle_neck:", "z_sample = torch.cat((z1, z2), dim=1)", uncomment=True)
    with contextlib.redirect_stdout(io.StringIO()):
        exec(compile(code, "pkl_file_models.py:302-332", "exec"), ns)
    out["dpca2_out"] = ns["z_sample"].numpy()
    out["dpca2_after"] = np.array([ns["self"].dimension_info["after_z1"], ns["self"].dimension_info["after_z2"]])
    ns = {"torch": torch, "data_pca": pm.data_pca, "self": types.SimpleNamespace(), "random_bottle_neck": True,
          "num_features": 4 * d, **{f"z{i + 1}": z.clone() for i, z in enumerate(zs)}}
    code = _reference_lines("pkl_file_models.py", 636, 679, "# if random_bottle_neck:", "z4 = torch.matmul(z4_p, v_4[:,ind_4].T) + mu_4", uncomment=True)
    torch.manual_seed(93)
    state = torch.random.get_rng_state()
    dim_p = torch.randint(8, int(4 * d / 2), (1,)).item()
    torch.random.set_rng_state(state)
    exec(compile(code, "pkl_file_models.py:636-679", "exec"), ns)
    assert ns["dim_p"] == dim_p
    out["dpca4_dim_p"] = np.array(dim_p)
    out["dpca4_out"] = torch.cat([ns[f"z{i + 1}"] for i in range(4)], 1).numpy()
    out["dpca4_after"] = np.array([ns["self"].dimension_info[f"after_z{i + 1}"] for i in range(4)])
    np.savez_compressed(os.path.join(HERE, "bottleneck.npz"), **out)


def gen_task_aware_tail():
    """Row a5-i: lines ``:312-364`` of ``task_aware()`` (``train_pkl_file_task_aware.py``) executed as written, with the
    discriminator passed in as a stub that returns seeded (logits, fmaps) of the shapes ``msstftd.py:185-192`` produces
    (3 sub-discriminators x 5 feature maps, scaled down) -- the wav2vec2 front half of the function cannot run here."""
    import torch.nn.functional as F
    code = _reference_lines("train_pkl_file_task_aware.py", 312, 364, "enhanced_length = enhanced_audio.shape[-1]", "discriminator_loss += F.mse_loss(y_gt")
    code = code.replace('device="cuda:0"', 'device="cpu"')
    g = torch.Generator().manual_seed(364)
    B, Le, Lc = 2, 958, 24000                                   # the reference's own comment: argmax sequence of 958 frames
    enhanced = torch.randint(0, 32, (B, Le), generator=g)       # torch.argmax(logits, -1)
    clean = torch.randn(B, Lc, generator=g) * 0.1
    shapes = [[(B, 4, 23, 17), (B, 4, 12, 17), (B, 4, 6, 17), (B, 4, 3, 17), (B, 4, 3, 17)],
              [(B, 4, 11, 33), (B, 4, 6, 33), (B, 4, 3, 33), (B, 4, 2, 33), (B, 4, 2, 33)],
              [(B, 4, 47, 9), (B, 4, 24, 9), (B, 4, 12, 9), (B, 4, 6, 9), (B, 4, 6, 9)]]
    lshapes = [(B, 1, 3, 17), (B, 1, 2, 33), (B, 1, 6, 9)]
    outs = {}

    def make(tag):
        fm = [[torch.randn(*s, generator=g) for s in sub] for sub in shapes]
        lg = [torch.randn(*s, generator=g) for s in lshapes]
        outs[tag] = (lg, fm)
        return lg, fm

    calls = []

    def disc(x):
        calls.append(tuple(x.shape))
        return make("enh" if len(calls) == 1 else "gt")

    ns = {"torch": torch, "np": np, "F": F, "disc": disc, "enhanced_audio": enhanced, "reconstructed_clean_waveforms": clean}
    exec(compile(code, "train_pkl_file_task_aware.py:312-364", "exec"), ns)
    assert calls == [(B, 1, Lc), (B, 1, Lc)]
    sv = {"enhanced": enhanced.numpy(), "clean": clean.numpy(), "mse": np.array(ns["mse_loss"].item()),
          "disc_loss": np.array(ns["discriminator_loss"].item())}
    for tag in ("enh", "gt"):
        lg, fm = outs[tag]
        for i, t in enumerate(lg):
            sv[f"{tag}_logit_{i}"] = t.numpy()
        for i, sub in enumerate(fm):
            for j, t in enumerate(sub):
                sv[f"{tag}_fmap_{i}_{j}"] = t.numpy()
    np.savez_compressed(os.path.join(HERE, "task_aware_tail.npz"), **sv)


if __name__ == "__main__":
    gen_stft()
    gen_dpca()
    gen_spec_losses()
    gen_var_topk()
    gen_istft()
    gen_cosine()
    gen_sgmse()
    gen_bottleneck()
    gen_task_aware_tail()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))
