"""ndtb_fitness_noalloc (include/ndt_b200/ndt_c_api.h): getFitnessScore with its scratch in a static device array, for callers that
cannot call cudaMalloc.  No counterpart in the reference.  Checked through the C ABI
against the existing entry point, ndtb_fitness, on two calls in a row that share the scratch array.  Named test_zz_* so that pytest
collects it last."""
import numpy as np
import pytest


def test_noalloc_entry_is_declared_and_exported():
    """CPU: the symbol loads; a size beyond the static array is rejected with a clear text before any device call."""
    from xchu_slam_b200 import build, ndt
    build.build_cuda()
    L = ndt.load_library()
    assert hasattr(L, "ndtb_fitness_noalloc") and hasattr(L, "ndtb_noalloc_last_error")
    with pytest.raises(ndt.NdtError, match="exceed the scratch array"):
        ndt.fitness_noalloc(0, 16, ndt.NOALLOC_MAX_SOURCE + 1, 16, 10, np.eye(4, dtype=np.float32), 1.0, 16)
    with pytest.raises(ndt.NdtError, match="null"):
        ndt.fitness_noalloc(0, 16, 4, 16, 4, np.eye(4, dtype=np.float32), 1.0, 0)


@pytest.mark.gpu
def test_noalloc_fitness_matches_existing_entry_on_two_calls_in_a_row():
    import torch
    from xchu_slam_b200 import ndt, synth
    src, tgt, gt, guess = synth.make_case(n_source=6000, n_target=80000, seed=synth.BASE_SEED + 21)
    g = ndt.NormalDistributionsTransform(device=0)
    g.setResolution(1.0)
    g.setTransformationEpsilon(0.01)
    g.setMaximumIterations(30)
    g.setInputTarget(tgt)
    g.setInputSource(src)
    g.align(guess)
    T_final = g.result()["T"]
    fmax = float(np.finfo(np.float64).max)
    # the existing entry point: (transform, max_range) -> score
    cases = [(T_final, fmax), (np.asarray(guess, dtype=np.float32), fmax), (T_final, 0.05), (np.asarray(guess, dtype=np.float32), 1e-12)]
    want = [g.getFitnessScore(mr, T=T) for T, mr in cases]
    assert want[0] < want[1] and want[3] == fmax          # the guess is farther than the result; nothing within 1e-12

    dev = torch.device("cuda", 0)
    s_dev = torch.from_numpy(np.ascontiguousarray(src[:, :4], dtype=np.float32)).to(dev)
    t_dev = torch.from_numpy(np.ascontiguousarray(tgt[:, :4], dtype=np.float32)).to(dev)
    out = torch.zeros(len(cases), dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    # (one call first: the driver loads a kernel's code on its first launch, which is not the entry point allocating)
    ndt.fitness_noalloc(0, s_dev.data_ptr(), 16, t_dev.data_ptr(), 16, T_final, fmax, out.data_ptr(), stream)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    # calls in a row on one stream, no synchronisation between them: every call re-arms the shared scratch.  The second call's
    # distances are LARGER than the first's: minima left over from the first call would show.
    for k, (T, mr) in enumerate(cases):
        ndt.fitness_noalloc(0, s_dev.data_ptr(), s_dev.shape[0], t_dev.data_ptr(), t_dev.shape[0], T, mr, out.data_ptr() + 8 * k, stream)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] == free0          # the calls took no device memory
    got = out.cpu().numpy()
    for k, (w, q) in enumerate(zip(want, got)):
        assert abs(q - w) <= 1e-9 * abs(w), (k, q, w)      # the tolerance of the other getFitnessScore tests

    # the same again, in the other order, after the scratch has been used: still the existing entry point's values
    out.zero_()
    for k in reversed(range(len(cases))):
        T, mr = cases[k]
        ndt.fitness_noalloc(0, s_dev.data_ptr(), s_dev.shape[0], t_dev.data_ptr(), t_dev.shape[0], T, mr, out.data_ptr() + 8 * k, stream)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    for k, (w, q) in enumerate(zip(want, got)):
        assert abs(q - w) <= 1e-9 * abs(w), (k, q, w)

    # edge cases: an empty source and an empty target give getFitnessScore's "nothing in range" value; too large is rejected
    ndt.fitness_noalloc(0, 0, 0, t_dev.data_ptr(), t_dev.shape[0], T_final, fmax, out.data_ptr(), stream)
    ndt.fitness_noalloc(0, s_dev.data_ptr(), s_dev.shape[0], 0, 0, T_final, fmax, out.data_ptr() + 8, stream)
    torch.cuda.synchronize()
    assert out[0].item() == fmax and out[1].item() == fmax
    with pytest.raises(ndt.NdtError, match="exceed the scratch array"):
        ndt.fitness_noalloc(0, s_dev.data_ptr(), ndt.NOALLOC_MAX_SOURCE + 1, t_dev.data_ptr(), t_dev.shape[0], T_final, fmax, out.data_ptr(), stream)
