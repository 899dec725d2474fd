"""The command line of the reference's `sartsolver` (source/arguments.cpp:86-236, time ranges :12-79) as re-implemented
in mpi-cuda-sartsolver_b200/host/arguments_b200.cpp: names, short forms, defaults, range checks, messages."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "mpi-cuda-sartsolver_b200", "host", "sart_b200_args")


def run(*argv):
    if not os.path.exists(EXE):
        subprocess.check_call(["make", "-C", os.path.dirname(EXE), "sart_b200_args"])
    out = subprocess.run([EXE, *argv], capture_output=True, text=True)
    return out.returncode, (json.loads(out.stdout) if out.returncode == 0 else out.stderr.strip())


def test_defaults_are_the_references():
    rc, c = run("rtm.h5", "image.h5")
    assert rc == 0
    assert c["input_files"] == ["rtm.h5", "image.h5"]
    assert c["output_file"] == "solution.h5" and c["raytransfer_name"] == "with_reflections"
    assert c["time_range"] == "" and c["laplacian_file"] == ""
    assert (c["wavelength_threshold"], c["ray_density_threshold"], c["ray_length_threshold"]) == (50.0, 1e-6, 1e-6)
    assert (c["conv_tolerance"], c["beta_laplace"], c["relaxation"]) == (1e-5, 2e-2, 1.0)
    assert (c["max_iterations"], c["max_cached_frames"], c["max_cached_solutions"]) == (2000, 100, 100)
    assert (c["logarithmic"], c["no_guess"], c["use_cpu"], c["parallel_read"]) == (0, 0, 0, 0)
    assert c["time_intervals"] == [[0, "inf", 0, 0]]


def test_every_option_short_and_long():
    rc, c = run("-o", "out.h5", "-t", "1:2", "-w", "30", "-d", "1e-4", "-r", "2e-4", "-m", "77", "-c", "1e-7", "-l", "lap.h5",
                "-b", "0.05", "-R", "0.7", "-n", "no_reflections", "-L", "--max_cached_frames", "5",
                "--max_cached_solutions=7", "--no_guess", "--use_cpu", "--parallel_read", "a.h5", "b.h5", "c.h5")
    assert rc == 0
    assert c["input_files"] == ["a.h5", "b.h5", "c.h5"]
    assert (c["output_file"], c["laplacian_file"], c["raytransfer_name"]) == ("out.h5", "lap.h5", "no_reflections")
    assert (c["wavelength_threshold"], c["ray_density_threshold"], c["ray_length_threshold"]) == (30.0, 1e-4, 2e-4)
    assert (c["max_iterations"], c["conv_tolerance"], c["beta_laplace"], c["relaxation"]) == (77, 1e-7, 0.05, 0.7)
    assert (c["max_cached_frames"], c["max_cached_solutions"]) == (5, 7)
    assert (c["logarithmic"], c["no_guess"], c["use_cpu"], c["parallel_read"]) == (1, 1, 1, 1)
    rc, c2 = run("--output_file", "x.h5", "--max_iterations=3", "--relaxation", "1", "a", "b")
    assert rc == 0 and (c2["output_file"], c2["max_iterations"]) == ("x.h5", 3)


@pytest.mark.parametrize("argv,fragment", [
    (["-d", "-1", "a", "b"], "ray_density_threshold must be >= 0"),
    (["-r", "-0.5", "a", "b"], "ray_length_threshold must be >= 0"),
    (["-m", "0", "a", "b"], "max_iterations must be >= 1"),
    (["-c", "0", "a", "b"], "conv_tolerance must be > 0"),
    (["-R", "0", "a", "b"], "relaxation must be within (0, 1] interval"),
    (["-R", "1.5", "a", "b"], "relaxation must be within (0, 1] interval"),
    (["-b", "-2", "a", "b"], "beta_laplace must be positive"),
    (["--max_cached_frames", "0", "a", "b"], "max_cached_frames must be positive"),
    (["--max_cached_solutions", "-3", "a", "b"], "max_cached_solutions must be positive"),
    (["only_one.h5"], "At least two input file"),
    ([], "At least two input file"),
    (["--bogus", "a", "b"], "Unknown argument"),
    (["-m"], "Too few arguments"),
    (["-m", "ten", "a", "b"], "Failed to parse"),
])
def test_rejections(argv, fragment):
    rc, err = run(*argv)
    assert rc == 1 and fragment in err


def test_time_ranges():
    rc, c = run("-t", "20.5:40.1, 45.2:51:5:0.05,", "a", "b")  # blank after the comma and a trailing comma are fine
    assert rc == 0 and c["time_intervals"] == [[20.5, 40.1, 0, 0], [45.2, 51, 5, 0.05]]
    for bad, fragment in [("5", None), ("3:2", "upper limit"), ("-1:2", "must be positive"), ("1:2:5", "Time step"),
                          ("1:5:2:3", "Synchronization threshold"), ("1:2:3:4:5", "Too many values"),
                          ("12,", "Unable to recognize"), ("a:b", "Unable to convert")]:
        rc, out = run("-t", bad, "a", "b")
        if fragment is None:
            # a one-character range is skipped by the reference's loop bound (source/arguments.cpp:25): no intervals
            assert rc == 0 and out["time_intervals"] == []
        else:
            assert rc == 1 and fragment in out, (bad, out)
