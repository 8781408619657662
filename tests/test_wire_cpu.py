"""Decoding of the allocator's wire format (AllocationRequest JSON, types.go:5-10 and the json
tags of trainingjob.go:17-66) — host logic only, no GPU."""
import json

import pytest

from vodascheduler_b200 import wire
from vodascheduler_b200.algorithm import Time

# what Go's encoding/json emits for an AllocationRequest with one job (field order of the structs)
BODY = """{"SchedulerID":"nvidia-gtx-1080ti","NumGpu":8,"AlgorithmName":"ElasticTiresias",
 "ReadyJobs":[{"job_name":"tf2-keras-mnist-elastic-20211031-135003","job_category":"tf2-keras-mnist-elastic",
   "user":"","priority":1,"gpu_type":"nvidia-gtx-1080ti","status":"Waiting",
   "submit_timestamp":"2021-10-31T13:50:03.123456789+08:00",
   "config":{"num_proc":2,"min_num_proc":1,"max_num_proc":4,"epochs":30},
   "time_metrics":{"first_start_timestamp":"2021-10-31T06:00:00Z","running_seconds":12.5},
   "info":{"job_name":"tf2-keras-mnist-elastic-20211031-135003","estimate_remainning_time_seconds":321.5,
           "speedup":{"0":0,"1":1,"2":1.8,"4":3.1},"efficiency":{"1":1}}}]}"""


def test_parse_time_rfc3339():
    assert wire.parse_time("1970-01-01T00:00:00Z") == Time(0, 0)
    assert wire.parse_time("1970-01-01T00:00:01.5Z") == Time(1, 500000000)
    assert wire.parse_time("2021-10-31T13:50:03.123456789+08:00") == Time(1635659403, 123456789)
    assert wire.parse_time("2021-10-31T05:50:03-01:30") == Time(1635659403 + 5400, 0)  # 07:20:03Z
    assert wire.parse_time("0001-01-01T00:00:00Z") == wire.ZERO_TIME  # Go's zero time.Time
    for bad in ("2021-10-31", "2021-10-31 13:50:03Z", "2021-10-31T13:50:03", ""):
        with pytest.raises(ValueError):
            wire.parse_time(bad)


def test_time_round_trip():
    for t in (Time(0, 0), Time(1635659403, 123456789), Time(-62135596800, 0), Time(253402300799, 999999999)):
        assert wire.parse_time(wire.format_time(t)) == t


def test_decode_request():
    req = wire.AllocationRequest.from_json(BODY)
    assert (req.SchedulerID, req.NumGpu, req.AlgorithmName) == ("nvidia-gtx-1080ti", 8, "ElasticTiresias")
    (j,) = req.ReadyJobs
    assert j.Name == "tf2-keras-mnist-elastic-20211031-135003" and j.Priority == 1
    assert j.SubmitTime == Time(1635659403, 123456789)
    assert (j.Config.NumProc, j.Config.MinNumProc, j.Config.MaxNumProc, j.Config.Epochs) == (2, 1, 4, 30)
    assert j.Metrics.FirstStartTime == Time(1635660000, 0)
    assert j.Info.EstimatedRemainningTimeSec == 321.5
    f32 = lambda x: float(__import__("numpy").float32(x))  # map[string]float32: the values are float32
    assert j.Info.Speedup == {"0": 0.0, "1": 1.0, "2": f32(1.8), "4": f32(3.1)}


def test_decode_defaults_like_encoding_json():
    """Missing keys leave Go zero values; `info` is omitempty so absent means a nil pointer."""
    req = wire.AllocationRequest.from_json('{"AlgorithmName":"FIFO","ReadyJobs":[{"job_name":"a"}]}')
    assert req.NumGpu == 0 and req.SchedulerID == ""
    (j,) = req.ReadyJobs
    assert j.Info is None and j.SubmitTime == wire.ZERO_TIME and j.Metrics.FirstStartTime == wire.ZERO_TIME
    assert (j.Config.NumProc, j.Config.MinNumProc, j.Config.MaxNumProc) == (0, 0, 0)
    assert wire.AllocationRequest.from_json('{"ReadyJobs":null}').ReadyJobs == []


def test_encode_decode_round_trip():
    req = wire.AllocationRequest.from_json(BODY)
    again = wire.AllocationRequest.from_json(req.to_json())
    assert json.loads(again.to_json()) == json.loads(req.to_json())
    assert again.ReadyJobs[0] == req.ReadyJobs[0]


def test_handler_rejects_without_touching_the_gpu():
    """Bad JSON and an unknown policy answer 400 before any engine call (resource_allocator.go:56-63,
    89-94), so these run without a GPU."""
    code, body = wire.handle_allocation("{not json")
    assert code == 400
    code, body = wire.handle_allocation('{"AlgorithmName":"Gandiva","NumGpu":4,"ReadyJobs":[]}')
    assert code == 400 and json.loads(body) == {"error": "Not found"}
    code, _ = wire.handle_allocation('{"AlgorithmName":"FIFO","ReadyJobs":[{"job_name":"a","submit_timestamp":"yesterday"}]}')
    assert code == 400


def test_float_fields_round_once_to_float32():
    """encoding/json parses a float32 field straight from the decimal text (ParseFloat(s, 32)).  A
    decimal a hair above the midpoint of two float32 values rounds UP there; going through float64
    first lands exactly on the midpoint and ties-to-even rounds it DOWN."""
    import numpy as np
    from vodascheduler_b200 import wire
    above = "1.000000059604644775390625000000000000000000001"  # 1 + 2^-24 + epsilon
    up = float(np.float32(1.0) + np.float32(2.0 ** -23))
    assert float(np.float32(float(above))) == 1.0          # the double rounding this avoids
    assert wire.parse_float32(above) == up
    assert wire.parse_float32("1.000000059604644775390625") == 1.0   # the exact midpoint: ties to even
    assert wire.parse_float32("0.1") == float(np.float32(0.1))
    body = '{"SchedulerID":"s","NumGpu":1,"AlgorithmName":"ElasticSRJF","ReadyJobs":[{"job_name":"a",' \
           '"info":{"estimate_remainning_time_seconds":%s,"speedup":{"1":%s}}}]}' % (above, above)
    req = wire.AllocationRequest.from_json(body)
    assert req.ReadyJobs[0].Info.EstimatedRemainningTimeSec == up and req.ReadyJobs[0].Info.Speedup["1"] == up
