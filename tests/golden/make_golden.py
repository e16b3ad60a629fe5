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


if __name__ == "__main__":
    gen_stft()
    gen_dpca()
    gen_spec_losses()
    gen_var_topk()
    gen_istft()
    gen_cosine()
    gen_sgmse()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))
