"""The pure-python HDF5 reader behind `walk_imitation(ref_path="*.hdf5")` (the reference reads its clip
datasets with h5py: vnl_ray/tasks/trajectory_loaders.py:33-35, 90-97, 155-157, 189-223)."""
import os

import numpy as np
import pytest

from flybody_b200 import clips as C
from flybody_b200 import hdf5_lite as H
from flybody_b200.compile_fly import load_model
from flybody_b200.model_pack import asset_path


def _scipy_matlab_dir():
    import scipy.io
    d = os.path.join(os.path.dirname(scipy.io.__file__), "matlab", "tests", "data")
    return d if os.path.exists(os.path.join(d, "testhdf5_7.4_GLNX86.mat")) else None


def test_reads_a_file_written_by_the_hdf5_library_itself():
    """scipy ships one MATLAB 7.3 file = HDF5 written by the HDF5 C library (512-byte user block, superblock 0, symbol-table
    group, version-1 object headers).  The same variable is stored next to it in MATLAB-5 format, which scipy reads: the values
    must agree bit for bit."""
    d = _scipy_matlab_dir()
    if d is None:
        pytest.skip("scipy's MATLAB test data is not installed")
    import scipy.io
    with H.File(os.path.join(d, "testhdf5_7.4_GLNX86.mat")) as f:
        assert f.keys() == ["testdouble"] and "testdouble" in f and "nope" not in f
        ds = f["testdouble"]
        assert ds.shape == (9, 1) and ds.dtype == np.float64       # MATLAB stores the 1 x 9 row column-major
        got = ds[()]
    ref = scipy.io.loadmat(os.path.join(d, "testdouble_7.4_GLNX86.mat"))["testdouble"]
    assert ref.shape == (1, 9) and np.array_equal(got.T, ref)
    assert np.array_equal(got[:, 0], np.linspace(0, 2 * np.pi, 9))


def test_round_trip_of_every_storage_form(tmp_path):
    """Contiguous, chunked, chunked + shuffle + deflate (ragged edge chunks, a chunk B-tree two levels deep), scalars, integer
    and fixed-length string data, groups big enough for several symbol-table nodes and a two-level group B-tree."""
    rng = np.random.RandomState(0)
    big = rng.normal(size=(130, 7, 3))
    many = rng.randint(-5, 5, size=(70, 70)).astype(np.int32)
    tree = {
        "a": {"x": big, "y": rng.normal(size=(5,)).astype(np.float32), "names": np.array([b"coxa_T1_left", b"femur", b""])},
        "scalar": np.float64(2e-3), "lengths": np.arange(11, dtype=np.int64), "u8": np.arange(7, dtype=np.uint8),
        "many": many, "empty": np.zeros((0, 4)), "vlen": np.array(["coxa_abduct_T1_left", "", "tarsus2_T3_right", "\u00e9"], dtype=object),
        "g": {f"{i:04d}": {"v": np.full((2, 2), float(i))} for i in range(300)},
    }
    p = str(tmp_path / "t.hdf5")
    H.write_file(p, tree, dataset_options={"x": dict(chunks=(16, 4, 3), deflate=4, shuffle=True), "many": dict(chunks=(1, 70)),
                                           "lengths": dict(chunks=(4,), deflate=1)}, leaf_k=2, internal_k=2)
    with H.File(p) as f:
        assert sorted(f.keys()) == sorted(tree)
        assert np.array_equal(f["a"]["x"][()], big) and np.array_equal(f["a/x"][3:9], big[3:9])
        assert f["a"]["y"].dtype == np.float32 and np.array_equal(np.asarray(f["a"]["y"]), tree["a"]["y"])
        assert [s.decode() for s in f["a"]["names"]] == ["coxa_T1_left", "femur", ""]
        assert [b.decode("utf-8") for b in f["vlen"]] == ["coxa_abduct_T1_left", "", "tarsus2_T3_right", "\u00e9"]   # global-heap strings
        assert f["scalar"][()] == 2e-3 and f["scalar"].shape == ()
        assert np.array_equal(f["lengths"][()], np.arange(11)) and f["lengths"][4] == 4 and len(f["lengths"]) == 11
        assert np.array_equal(f["u8"][()], np.arange(7)) and f["u8"].dtype == np.uint8
        assert np.array_equal(f["many"][()], many) and f["empty"].shape == (0, 4)
        assert len(f["g"]) == 300 and f["g"].keys() == sorted(tree["g"])
        assert all(f["g"][k]["v"][()][0, 0] == float(int(k)) for k in ("0000", "0137", "0299"))
        with pytest.raises(KeyError):
            f["g"]["0300"]
    # default node sizes (what the library writes) as well
    H.write_file(p, {"g": tree["g"]})
    with H.File(p) as f:
        assert len(f["g"]) == 300 and f["g/0042/v"][()][1, 1] == 42.0


def test_unsupported_files_are_refused_by_name(tmp_path):
    p = str(tmp_path / "v2.h5")
    open(p, "wb").write(H.SIGNATURE + bytes([2, 8, 8, 0]) + bytes(64))
    with pytest.raises(H.Hdf5Unsupported, match="superblock version 2"):
        H.File(p)
    open(p, "wb").write(b"not an hdf5 file" * 8)
    with pytest.raises(H.Hdf5FormatError):
        H.File(p)


def test_walking_dataset_in_the_reference_layout_loads_as_the_clip_table(tmp_path):
    """Synthetic clips exported in the reference's dataset layout come back as the table `pack_clips` builds from them directly:
    numeric order of the zero-padded keys, [root_qpos | qpos] concatenation, joint columns matched by `id2name` (here shuffled on
    disk), xy origin at the first frame, timestep check."""
    M = load_model(asset_path("walk_imitation"))
    cl = [C.synth_walk_clip(M, T=70 + 3 * i, seed=i) for i in range(12)]
    p = str(tmp_path / "walk.hdf5")
    C.export_hdf5(p, M, cl, dataset_options={"joint_quat": dict(chunks=(32, 66, 4), deflate=3, shuffle=True)})
    want = C.pack_clips(M, cl)
    got = C.load_hdf5_clips(M, p)
    assert all(np.array_equal(a, b) for a, b in zip(got, want))
    with H.File(p) as f:
        assert len(f["trajectories"]) == 12 and f["trajectories"].keys()[0] == "00"
        assert np.array_equal(f["trajectory_lengths"][()], [c["qpos"].shape[0] for c in cl])
        assert [s.decode() for s in f["id2name"]["sites"]][0] == "claw_T1_left"
        # the reference's loader arithmetic on this file (trajectory_loaders.py:189-215)
        sn = f["trajectories"]["03"]
        qpos = np.concatenate((sn["root_qpos"][5:20], sn["qpos"][5:20]), axis=1)
        assert np.array_equal(qpos, cl[3]["qpos"][5:20])
    # a dataset whose joints are stored in another order and displaced in xy: matched by name, re-origined
    perm = np.random.RandomState(1).permutation(int(M["n_mocap_joints"]))
    from flybody_b200.hdf5_lite import write_file
    names = [M["_names_jnt"][int(j)].split("/", 1)[1].encode() for j in M["mocap_joint"]]
    c = cl[0]
    rq = c["qpos"][:, :7].copy(); rq[:, :2] += [3.0, -2.0]
    write_file(p, {"trajectories": {"0": dict(root_qpos=rq, qpos=c["qpos"][:, 7:][:, perm], root_qvel=c["qvel"][:, :6],
                                              qvel=c["qvel"][:, 6:][:, perm], root2site=c["root2site"], joint_quat=c["joint_quat"][:, perm])},
                   "timestep_seconds": np.float64(M["control_dt"]), "id2name": {"joints": np.array([names[i] for i in perm])}})
    rows, off, ln = C.load_hdf5_clips(M, p)
    w0 = C.pack_clips(M, [c])[0]
    assert np.abs(rows - w0).max() < 1e-6     # xy re-origin of the displaced copy rounds differently in the last float32 bit
    write_file(p, {"trajectories": {"0": dict(root_qpos=rq, qpos=c["qpos"][:, 7:], root_qvel=c["qvel"][:, :6], qvel=c["qvel"][:, 6:],
                                              root2site=c["root2site"], joint_quat=c["joint_quat"])}, "timestep_seconds": np.float64(0.01)})
    with pytest.raises(ValueError, match="timestep"):
        C.load_hdf5_clips(M, p)


def test_flight_dataset_in_the_reference_layout(tmp_path):
    M = load_model(asset_path("flight_imitation"))
    cl = [C.synth_flight_clip(T=60, seed=i) for i in range(3)]
    p = str(tmp_path / "flight.hdf5")
    C.export_hdf5(p, M, cl, kind="flight")
    rows, off, ln = C.load_hdf5_clips(M, p)
    want = C.pack_clips(M, cl)
    assert rows.shape == want[0].shape and np.abs(rows - want[0]).max() < 1e-6 and np.array_equal(ln, want[2])
    with H.File(p) as f:
        assert f["trajectories"]["1"]["com_qpos"].shape == (60, 7) and "id2name" not in f
