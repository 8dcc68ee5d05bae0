#!/usr/bin/env python
"""Generate the golden fixtures in this directory FROM THE REFERENCE'S OWN TESTS.

Run in the build container only (needs /root/reference; the GPU box has no copy):

    python tests/golden/make_golden.py

Nothing from the reference is copied into the repo.  This script
  * parses the reference's unit-test files with ``ast``,
  * executes *their* numpy oracles (``adamw_step`` / ``adam_step``) on *their*
    inputs (OpTest classes seed ``np.random.seed(123)`` in ``setUpClass``,
    python/paddle/fluid/tests/unittests/eager_op_test.py:327-339, and draw the
    inputs in ``setUp`` in a fixed order, reproduced below call for call),
  * reads the ten literal bucket-assignment cases out of the AST of
    test_imperative_group.py,
and stores inputs + expected outputs as small ``.npz`` / ``.json`` fixtures.  The
oracle (oracle/) is then pinned against these fixtures by tests/test_oracle_golden.py
with the tolerances the reference's tests use.
"""
import ast
import json
import os
import sys

import numpy as np

REF = "/root/reference/python/paddle/fluid/tests/unittests"
OUT = os.path.dirname(os.path.abspath(__file__))


def extract_function(path, name):
    """Return the reference's function ``name`` compiled from its own source."""
    src = open(path).read()
    tree = ast.parse(src)
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            mod = ast.Module(body=[node], type_ignores=[])
            ns = {"np": np}
            exec(compile(mod, path, "exec"), ns)
            return ns[name]
    raise KeyError(name)


def draw_adam_inputs(shape, lr, beta1, beta2):
    """Input draws of TestAdamW/TestAdamOp* setUp, in the reference's call order."""
    param = np.random.uniform(-1, 1, shape).astype("float32")
    grad = np.random.uniform(-1, 1, shape).astype("float32")
    moment1 = np.random.uniform(-1, 1, shape).astype("float32")
    moment2 = np.random.random(shape).astype("float32")
    return {
        "Param": param,
        "Grad": grad,
        "Moment1": moment1,
        "Moment2": moment2,
        "LearningRate": np.array([lr]).astype("float32"),
        "Beta1Pow": np.array([beta1**10]).astype("float32"),
        "Beta2Pow": np.array([beta2**10]).astype("float32"),
    }


def save_case(store, prefix, inputs, attrs, outs):
    for k, v in inputs.items():
        store[f"{prefix}/in/{k}"] = np.asarray(v)
    for k, v in attrs.items():
        store[f"{prefix}/attr/{k}"] = np.asarray(v)
    for k, v in zip(("ParamOut", "Moment1Out", "Moment2Out"), outs):
        store[f"{prefix}/out/{k}"] = np.asarray(v)


def gen_adamw():
    # test_adamw_op.py:27-60 (adamw_step), :104-211 (TestAdamW, TestAdamW2)
    adamw_step = extract_function(f"{REF}/test_adamw_op.py", "adamw_step")
    store = {}
    np.random.seed(123)
    inputs = draw_adam_inputs((102, 105), 0.004, 0.78, 0.836)
    attrs = {"epsilon": 1e-4, "beta1": 0.78, "beta2": 0.836, "coeff": 0.5, "with_decay": True}
    save_case(store, "TestAdamW", inputs, attrs, adamw_step(inputs, attrs))
    np.random.seed(123)
    inputs = draw_adam_inputs((2, 2), 0.004, 0.78, 0.836)
    attrs = {"epsilon": 1e-4, "beta1": 0.78, "beta2": 0.836, "lr_ratio": 0.1, "coeff": 0.5, "with_decay": True}
    save_case(store, "TestAdamW2", inputs, attrs, adamw_step(inputs, attrs))
    np.savez_compressed(f"{OUT}/adamw_optest.npz", **store)


def gen_adam():
    # test_adam_op.py:231-262 (adam_step), :61-227 (TestAdamOp1/2, OnlyTail, MultipleSteps)
    adam_step = extract_function(f"{REF}/test_adam_op.py", "adam_step")
    store = {}
    np.random.seed(123)
    inputs = draw_adam_inputs((102, 105), 0.004, 0.78, 0.836)
    attrs = {"epsilon": 1e-4, "beta1": 0.78, "beta2": 0.836}
    save_case(store, "TestAdamOp1", inputs, attrs, adam_step(inputs, attrs))
    np.random.seed(123)
    inputs = draw_adam_inputs((102, 105), 0.001, 0.9, 0.999)
    attrs = {"epsilon": 1e-8, "beta1": 0.9, "beta2": 0.999}
    save_case(store, "TestAdamOp2", inputs, attrs, adam_step(inputs, attrs))
    np.random.seed(123)
    inputs = draw_adam_inputs(3, 0.001, 0.9, 0.999)
    save_case(store, "TestAdamOnlyTailOp", inputs, attrs, adam_step(inputs, attrs))
    # TestAdamOpMultipleSteps: 10 steps, outputs feed back, fresh grad per step (:197-227)
    np.random.seed(123)
    inputs = draw_adam_inputs((102, 105), 0.001, 0.9, 0.999)
    for k, v in inputs.items():
        store[f"MultipleSteps/in/{k}"] = v.copy()
    for k, v in attrs.items():
        store[f"MultipleSteps/attr/{k}"] = np.asarray(v)
    for step in range(10):
        store[f"MultipleSteps/step{step}/Grad"] = inputs["Grad"].copy()
        p, m1, m2 = adam_step(inputs, attrs)
        b1p = inputs["Beta1Pow"] * 0.9
        b2p = inputs["Beta2Pow"] * 0.999
        if step in (0, 9):  # keep the fixture small: first and last step only
            store[f"MultipleSteps/step{step}/ParamOut"] = p
            store[f"MultipleSteps/step{step}/Moment1Out"] = m1
            store[f"MultipleSteps/step{step}/Moment2Out"] = m2
        store[f"MultipleSteps/step{step}/Beta1PowOut"] = b1p
        store[f"MultipleSteps/step{step}/Beta2PowOut"] = b2p
        inputs["Param"], inputs["Moment1"], inputs["Moment2"] = p, m1, m2
        inputs["Beta1Pow"], inputs["Beta2Pow"] = b1p, b2p
        inputs["Grad"] = np.random.uniform(-1, 1, (102, 105)).astype("float32")
    np.savez_compressed(f"{OUT}/adam_optest.npz", **store)


def gen_check_finite():
    # test_amp_check_finite_and_scale_op.py:29-90: x (1024,1024) random, scale random(1);
    # expected Out = x / scale, FoundInfinite 0 / 1 (NaN at [128][128]) / 1 (Inf).
    # The full tensors are regenerated from the seed by the test; the fixture keeps the
    # scale, a strided sample of x and of the expected output, and the flags.
    np.random.seed(123)
    x = np.random.random((1024, 1024)).astype(np.float32)
    scale = np.random.random((1)).astype(np.float32)
    out = x / scale
    np.savez_compressed(
        f"{OUT}/check_finite_optest.npz",
        scale=scale,
        x_sample=x[::16, ::16].copy(),
        out_sample=out[::16, ::16].copy(),
        x_sum=np.array([x.astype(np.float64).sum()]),
        found_clean=np.array([0]),
        found_nan=np.array([1]),
        found_inf=np.array([1]),
        poison_index=np.array([128, 128]),
    )


def gen_squared_l2_norm():
    # test_squared_l2_norm_op.py:72-84 (TestL2LossOp) and :33-70 (fp16-vs-fp32 cases)
    from numpy import linalg as LA

    store = {}
    np.random.seed(123)
    X = np.random.uniform(-1, 1, (13, 19)).astype("float32")
    X[np.abs(X) < 0.05] = 0.1
    store["TestL2LossOp/X"] = X
    store["TestL2LossOp/Out"] = np.array([np.square(LA.norm(X))])
    for name, lo, hi, shape in (
        ("F16Op", -0.1, 0.1, (8, 5, 10)),
        ("F16Op1", -2.0, 2.0, (30, 10)),
        ("F16Op2", -5.0, 5.0, (20, 10, 20)),
    ):
        # unittest.TestCase, not OpTest: unseeded in the reference; seeded here.
        np.random.seed(123)
        store[f"{name}/X"] = np.random.uniform(lo, hi, shape).astype("float32")
    np.savez_compressed(f"{OUT}/squared_l2_norm_optest.npz", **store)


def gen_bucket_cases():
    """Literal cases of test_imperative_group.py:30-172, read from its AST."""
    path = f"{REF}/test_imperative_group.py"
    tree = ast.parse(open(path).read())
    cases = []
    for cls in tree.body:
        if not (isinstance(cls, ast.ClassDef) and cls.name == "TestDataParallelGroup"):
            continue
        for fn in cls.body:
            if not (isinstance(fn, ast.FunctionDef) and fn.name.startswith("test_construct_group")):
                continue
            tensors, call_args, expected = [], None, None
            for node in ast.walk(fn):
                if isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute):
                    if node.func.attr == "create_varbase":
                        dtype = ast.literal_eval(node.args[0])
                        shape = ast.literal_eval(node.args[1])
                        tensors.append((node.lineno, dtype, shape))
                    elif node.func.attr == "assign_group_by_size":
                        call_args = [ast.literal_eval(a) for a in node.args[1:]]
                    elif node.func.attr == "assertEqual":
                        expected = ast.literal_eval(node.args[0])
            tensors.sort()
            case = {
                "name": fn.name,
                "source": f"test_imperative_group.py:{fn.lineno}",
                "dtypes": [t[1] for t in tensors],
                "shapes": [t[2] for t in tensors],
                "is_sparse": call_args[0],
                "limits": call_args[1],
                "tensor_indices": call_args[2] if len(call_args) > 2 else [],
                "expected": expected,
            }
            cases.append(case)
    cases.sort(key=lambda c: c["name"])
    assert len(cases) == 10, len(cases)
    with open(f"{OUT}/bucket_cases.json", "w") as f:
        json.dump(cases, f, indent=1)


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("reference tree not present; fixtures can only be regenerated in the build container")
    gen_adamw()
    gen_adam()
    gen_check_finite()
    gen_squared_l2_norm()
    gen_bucket_cases()
    print("fixtures written to", OUT)
