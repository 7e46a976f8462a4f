"""Row N2: the per-item input pipeline of TalkingFaceFrameDataset (dataset.py:78-136).
CPU: the oracle restatement against the golden fixture made from the reference's call sites with the reference's own
libraries (oracle/make_golden_audio.py), index arithmetic against Python floats over a sweep, host tables.
GPU: b2d_audio_window / b2d_frame_gather through the C ABI against the same fixture -- indices and frames bit-exact,
audio within the fp32 bar (1e-4)."""
import numpy as np
import pytest
import torch

from oracle import audio_window as A

TOL_FP32 = 1e-4       # north_star: relative error <= 1e-4 in fp32


def _store(g, device):
    import lipreading_video_generation_b200 as P
    st = P.DeviceVideoStore(device)
    for v in g["videos"]:
        st.add(v["name"], v["frames"], v["wave"], v["fps"], v["sr"])
    return st


def test_oracle_restatement_matches_reference_call_sites(golden):
    g = golden("audio_window.pt")
    vids = {v["name"]: v for v in g["videos"]}
    for k, (name, fe) in enumerate(g["items"]):
        v = vids[name]
        assert A.window_indices(fe, v["T"], v["fps"], v["sr"]) == tuple(g["indices"][k].tolist())
        o = A.audio_window(v["wave"][0].numpy(), fe, v["T"], v["fps"], v["sr"])
        ref = g["input_values"][k].numpy()
        assert np.abs(o - ref).max() <= TOL_FP32
        assert np.linalg.norm(o - ref) / np.linalg.norm(ref) <= TOL_FP32
        assert torch.equal(A.frame_to_tensor(v["frames"][g["indices"][k, 0]].numpy()), g["output_frame"][k])
        assert torch.equal(A.frame_to_tensor(v["frames"][0].numpy()), g["input_frame"][k])


def test_sinc_table_and_biquad_coefficients_equal_torchaudio():
    torchaudio = pytest.importorskip("torchaudio")
    from lipreading_video_generation_b200 import dataset as D
    rs = torchaudio.transforms.Resample(orig_freq=1, new_freq=16000)          # dataset.py:53-55 with size(0) == 1
    assert torch.equal(D.sinc_rows_1hz_to_16k(), rs.kernel[:4000, 0, :])
    assert torch.equal(A.sinc_kernel_rows()[0], rs.kernel[:4000, 0, :])
    for sr in (16000, 44100, 48000):
        x = torch.zeros(1, 8); x[0, 0] = 0.25                                  # impulse response pins b0, b1, b2, a1, a2
        y = torchaudio.functional.highpass_biquad(x, sr, 300)[0]
        b0, b1, b2, a1, a2 = D.highpass_biquad_coeffs(sr)
        h = [0.25 * b0, 0.25 * b1 - a1 * (0.25 * b0)]
        assert abs(float(y[0]) - h[0]) <= 1e-7 and abs(float(y[1]) - h[1]) <= 1e-6
        assert D.highpass_biquad_coeffs(sr) == [*A.highpass_coeffs(sr)[0], *A.highpass_coeffs(sr)[1]]


def test_host_tables_and_store_layout(golden):
    import lipreading_video_generation_b200 as P
    g = golden("audio_window.pt")
    st = _store(g, "cpu").finalize()
    assert st.frames.shape[0] == sum(v["T"] for v in g["videos"]) and st.frames.dtype == torch.uint8
    assert st.first_frame.tolist() == [0, 40, 115, 145] and st.wave_len.tolist() == [v["wave"].numel() for v in g["videos"]]
    items = [P.FrameItem(n, 0, fe) for n, fe in g["items"]]
    b = P.TalkingFaceFrameBatcher(st, items)
    vid, fend, tot, fps, sr, bq = b.host_tables(range(len(items)))
    assert vid.tolist() == [0, 0, 0, 1, 1, 2, 2, 3, 3, 1] and fend.tolist() == [fe for _, fe in g["items"]]
    assert tot.tolist() == [40, 40, 40, 75, 75, 30, 30, 50, 50, 75] and fps.dtype == np.float64 and bq.shape == (10, 5)
    with pytest.raises(RuntimeError):
        b.batch([0])                                     # CPU store: no fallback
    with pytest.raises(ValueError):
        P.DeviceVideoStore("cpu").add("x", torch.zeros(2, 4, 4, 3, dtype=torch.uint8), torch.zeros(1, 10), 0, 16000)


@pytest.mark.gpu
def test_device_batch_matches_reference_fixture(golden):
    import lipreading_video_generation_b200 as P
    g = golden("audio_window.pt")
    st = _store(g, "cuda")
    items = [P.FrameItem(n, 0, fe) for n, fe in g["items"]]
    b = P.TalkingFaceFrameBatcher(st, items)
    fin, fout, audio = b.batch(list(range(len(items))))
    assert torch.equal(b.last_indices.cpu(), g["indices"])                    # bit-exact frame / sample indices
    assert torch.equal(fin.cpu(), g["input_frame"]) and torch.equal(fout.cpu(), g["output_frame"])
    iv = audio["input_values"]
    assert iv.shape == (len(items), 1, 4000)                                  # the collated (B, 1, 4000) of train.py:119
    ref = g["input_values"]
    assert (iv[:, 0].cpu() - ref).abs().max().item() <= TOL_FP32
    assert ((iv[:, 0].cpu() - ref).norm() / ref.norm()).item() <= TOL_FP32
    b.check()
    sub = b.batch([7, 2])                                                     # ragged order / subset
    assert torch.equal(b.last_indices.cpu(), g["indices"][[7, 2]])
    assert (sub[2]["input_values"][:, 0].cpu() - ref[[7, 2]]).abs().max().item() <= TOL_FP32


@pytest.mark.gpu
def test_device_index_arithmetic_is_bit_exact_over_a_sweep():
    """Size-independent property at scale: 20000 random (frame_end, total_frames, fps, sr) against Python floats."""
    import lipreading_video_generation_b200 as P
    rng = np.random.default_rng(0)
    n = 20000
    fps = rng.choice([23.976, 24.0, 25.0, 29.97, 30.0, 50.0, 59.94, 12.5, 1000 / 33.0], n)
    sr = rng.choice([16000, 22050, 44100, 48000], n)
    tot = rng.integers(1, 2000, n)
    fend = rng.integers(0, 2200, n)
    st = P.DeviceVideoStore("cuda")
    st.add("v", torch.zeros(1, 2, 2, 3, dtype=torch.uint8), torch.zeros(1, 4), 25.0, 16000)
    b = P.TalkingFaceFrameBatcher(st, [])
    dev = lambda a: torch.from_numpy(a).cuda()
    vid = torch.zeros(n, dtype=torch.int32, device="cuda")
    out = torch.empty(n, 4000, device="cuda"); idx = torch.empty(n, 3, dtype=torch.int64, device="cuda")
    bq = torch.zeros(n, 5, device="cuda")
    from lipreading_video_generation_b200 import _lib as L
    d_fend, d_tot, d_fps, d_sr = dev(fend.astype(np.int64)), dev(tot.astype(np.int64)), dev(fps.astype(np.float64)), dev(sr.astype(np.int32))
    L.check(L.lib().b2d_audio_window(st.waves.data_ptr(), st.wave_offset.data_ptr(), st.wave_len.data_ptr(), vid.data_ptr(),
                                     d_fend.data_ptr(), d_tot.data_ptr(), d_fps.data_ptr(), d_sr.data_ptr(),
                                     b.sinc_rows.data_ptr(), bq.data_ptr(), n, out.data_ptr(), idx.data_ptr(),
                                     b.err_flag.data_ptr(), L.stream_ptr()), "b2d_audio_window")
    want = [A.window_indices(int(fend[i]), int(tot[i]), float(fps[i]), int(sr[i])) for i in range(n)]
    assert idx.cpu().tolist() == [list(w) for w in want]
    with pytest.raises(ValueError):
        b.check()                                        # 4-sample track: every window is empty -> flagged, NaN written
    assert torch.isnan(out).all()


@pytest.mark.gpu
def test_empty_batch_and_window_clipped_at_the_end_of_a_short_track(golden):
    """Edge cases: an empty index list launches nothing; a window that runs past the end of the track is clipped like
    the reference's Python slice (and still normalised over what is left)."""
    import lipreading_video_generation_b200 as P
    from lipreading_video_generation_b200 import _lib as L
    g = golden("audio_window.pt")
    st = _store(g, "cuda")
    b = P.TalkingFaceFrameBatcher(st, [P.FrameItem(n, 0, fe) for n, fe in g["items"]])
    l0 = L.lib().b2d_launch_count()
    fin, fout, audio = b.batch([])
    assert L.lib().b2d_launch_count() == l0
    assert fin.shape[0] == 0 and fout.shape[0] == 0 and audio["input_values"].shape == (0, 1, 4000)
    # a 1.0 s track at 16 kHz with 40 frames at 25 fps: frame 39 ends at 1.56 s -> window [21760, 24960) clipped to 16000
    st2 = P.DeviceVideoStore("cuda")
    wave = g["videos"][0]["wave"][:, :23000]
    st2.add("short.mp4", g["videos"][0]["frames"], wave, 25.0, 16000)
    b2 = P.TalkingFaceFrameBatcher(st2, [P.FrameItem("short.mp4", 0, 39)])
    _, _, a2 = b2.batch([0])
    assert b2.last_indices.cpu().tolist() == [[39, 21760, 24960]]          # indices are NOT clipped, the slice is
    want = A.audio_window(wave[0].numpy(), 39, 40, 25.0, 16000)            # numpy slicing clips the same way
    assert np.abs(a2["input_values"][0, 0].cpu().numpy() - want).max() <= TOL_FP32
    b2.check()
