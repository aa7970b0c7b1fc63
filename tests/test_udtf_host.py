"""Host-side behaviour of the UDTF mirror that needs no GPU: argument checks happen before any
device work, strings are packed the way the C-ABI expects, and the error classes map the way a
Hive integrator would rely on."""
import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import udtf


def test_pack_strings_offsets_and_bytes():
    data, off = udtf._pack_strings(["a", "", "héllo", b"\x00\x01"])
    assert off.tolist() == [0, 1, 1, 7, 9] and off.dtype == np.int64
    assert bytes(data) == b"a" + "héllo".encode() + b"\x00\x01"
    data, off = udtf._pack_strings([])
    assert len(data) == 0 and off.tolist() == [0]


def test_error_classes():
    assert issubclass(hm.UDFArgumentError, hm.HivemallError) and issubclass(hm.UDFArgumentError, ValueError)
    e = hm.UDFArgumentError(1, "bad option")
    assert e.status == 1 and "HM_ERR_INVALID_ARGUMENT" in str(e)


def test_ptr_helper_accepts_arrays_ints_and_none():
    a = np.zeros(4, np.float32)
    assert udtf._ptr(a).value == a.ctypes.data
    assert udtf._ptr(None).value in (None, 0) and udtf._ptr(1234).value == 1234
    with pytest.raises(TypeError):
        udtf._ptr("nope")


def test_bad_options_fail_before_any_device_work():
    # the option string is parsed first: these raise the argument error even without a GPU
    for fn, opts in [(hm.train_classifier, "-opt nope"), (hm.train_regressor, "-loss logloss"), (hm.train_fm, "-factors 0")]:
        with pytest.raises(hm.UDFArgumentError):
            fn(opts)


def test_mix_module_imports_without_distributed_init():
    from hivemall_b200 import mix
    assert mix.shard_rows(10, 0, 1) == (0, 10)
    mix.average_tensors([], world=1)        # world 1: nothing to do, no process group needed
