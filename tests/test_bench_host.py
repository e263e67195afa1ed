"""Host-side logic of bench.py that needs no GPU: the e2e copy-schedule selection."""

import bench


class _Pipe:
    serial = False


def _timed_for(pipe, ms):
    seen = []

    def timed(step, steps, warmup):
        seen.append(pipe.serial)
        return ms["one_stream" if pipe.serial else "overlapped"]

    return timed, seen


def test_measure_e2e_reports_the_faster_schedule_and_both_times(monkeypatch):
    monkeypatch.delenv("MVB200_E2E_SERIAL", raising=False)
    pipe = _Pipe()
    timed, seen = _timed_for(pipe, {"overlapped": 8.9, "one_stream": 11.0})
    ms, info = bench.measure_e2e(pipe, None, timed, 3, 1)
    assert seen == [False, True]
    assert ms == 8.9 and info["schedule"] == "overlapped"
    assert info["ms_per_step_by_schedule"] == {"overlapped": 8.9, "one_stream": 11.0}
    timed, _ = _timed_for(pipe, {"overlapped": 7.0, "one_stream": 6.1})
    ms, info = bench.measure_e2e(pipe, None, timed, 3, 1)
    assert ms == 6.1 and info["schedule"] == "one_stream" and "one stream" in info["api"]


def test_measure_e2e_can_be_forced(monkeypatch):
    pipe = _Pipe()
    for flag, name in (("1", "one_stream"), ("0", "overlapped")):
        monkeypatch.setenv("MVB200_E2E_SERIAL", flag)
        timed, seen = _timed_for(pipe, {"overlapped": 1.0, "one_stream": 2.0})
        _, info = bench.measure_e2e(pipe, None, timed, 3, 1)
        assert info["schedule"] == name and len(seen) == 1
