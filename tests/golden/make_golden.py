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


def gen_fused_adam():
    """TestFusedAdamOp (test_fused_adam_op.py:88-175) with its numpy oracle fused_adam_step (:23-85).

    The reference builds its six input lists as ``[[0] * num] * 6`` -- six references to ONE list -- so every
    assignment in its setUp loop lands in the same slot and all six "lists" end up holding the last value written,
    ``np.array([beta2_pow])``.  The draws still consume the generator.  Reproduced here statement for statement, so
    the fixture holds exactly the tensors that test feeds the operator.

    That accident is also why the reference's test passes: fused_adam_step adds ``epsilon`` to ``sqrt(m2)`` unscaled,
    while the kernel (fused_adam_kernel.cu:213-227, like adam_step in test_adam_op.py) uses
    ``epsilon * sqrt(1 - beta2_pow)``; on (102, 105) uniform data the two differ by up to 2e-4 where m2 is tiny,
    far outside the OpTest tolerance, on the one-element tensors the test really feeds they agree.  The oracle
    follows the kernel, so only the inputs the reference really tests are pinned here.
    """
    fused_adam_step = extract_function(f"{REF}/test_fused_adam_op.py", "fused_adam_step")
    np.random.seed(123)
    num = 10
    inputs_list = [[0] * num] * 6
    learning_rate, beta1, beta2, epsilon = 0.004, 0.78, 0.836, 1e-4
    beta1_pow, beta2_pow = beta1**10, beta2**10
    attrs = {"epsilon": epsilon, "beta1": beta1, "beta2": beta2, "chunk_size": 32 * 2048}
    for i in range(num):
        inputs_list[0][i] = np.random.uniform(-1, 1, (102, 105)).astype("float32")
        inputs_list[1][i] = np.random.uniform(-1, 1, (102, 105)).astype("float32")
        inputs_list[2][i] = np.random.uniform(-1, 1, (102, 105)).astype("float32")
        inputs_list[3][i] = np.random.random((102, 105)).astype("float32")
        inputs_list[4][i] = np.array([beta1_pow]).astype("float32")
        inputs_list[5][i] = np.array([beta2_pow]).astype("float32")
    names = ("Params", "Grads", "Moments1", "Moments2", "Beta1Pows", "Beta2Pows")
    inputs = {n: [(n.lower() + str(i), inputs_list[k][i]) for i in range(num)] for k, n in enumerate(names)}
    inputs["LearningRate"] = np.array([learning_rate]).astype("float32")
    outs = fused_adam_step(inputs, attrs, num)
    store = {"num": np.array([num]), "LearningRate": inputs["LearningRate"]}
    for k, v in attrs.items():
        store[f"attr/{k}"] = np.asarray(v)
    for n in names:
        for i in range(num):
            store[f"in/{n}/{i}"] = np.asarray(inputs[n][i][1])
    for n, arrs in zip(("ParamsOut", "Moments1Out", "Moments2Out", "Beta1PowsOut", "Beta2PowsOut"), outs):
        for i in range(num):
            store[f"out/{n}/{i}"] = np.asarray(arrs[i])
    np.savez_compressed(f"{OUT}/fused_adam_optest.npz", **store)


def gen_adamw_layerwise():
    """TestAdamWOpLayerwiseLR.test_adamw_op_dygraph (test_adamw_op.py:628-755): five AdamW steps over two Linear
    layers with lr_ratio = simple_lr_setting (:614-622), checked there with rtol 1e-6 against adamw_step.  The
    reference's gradients come from its autograd; here they are drawn (seed 2022 as in its setUp), everything
    else -- parameter names, shapes, hyper-parameters, the per-step beta powers, the two reference functions --
    is the reference's."""
    from functools import partial
    from types import SimpleNamespace

    adamw_step = extract_function(f"{REF}/test_adamw_op.py", "adamw_step")
    simple_lr_setting = extract_function(f"{REF}/test_adamw_op.py", "simple_lr_setting")
    simple_lr_fun = partial(simple_lr_setting, decay_rate=0.8, n_layers=2)
    np.random.seed(2022)
    learning_rate, weight_decay, beta1, beta2 = 0.001, 0.01, 0.9, 0.999
    shapes = {"linear_1.w_0": (13, 8), "linear_1.b_0": (8,), "linear_2.w_0": (8, 5), "linear_2.b_0": (5,)}
    store = {"steps": np.array([5])}
    state = {}
    for name, shape in shapes.items():
        w = np.random.uniform(-0.5, 0.5, shape).astype("float32") if name.endswith("w_0") else np.ones(shape, "float32")
        state[name] = [w, np.zeros_like(w), np.zeros_like(w)]
        store[f"{name}/lr_ratio"] = np.array([simple_lr_fun(SimpleNamespace(name=name))])
        store[f"{name}/param0"] = w.copy()
    for t in range(1, 6):
        for name, shape in shapes.items():
            grad = np.random.uniform(-1, 1, shape).astype("float32")
            param, m1, m2 = state[name]
            np_inputs = {"Param": param, "Grad": grad, "Moment1": m1, "Moment2": m2,
                         "LearningRate": np.array([learning_rate]).astype("float32"),
                         "Beta1Pow": np.array([beta1**t]).astype("float32"),
                         "Beta2Pow": np.array([beta2**t]).astype("float32")}
            np_attrs = {"epsilon": 1e-8, "beta1": beta1, "beta2": beta2,
                        "lr_ratio": float(store[f"{name}/lr_ratio"][0]), "coeff": weight_decay, "with_decay": True}
            state[name] = list(adamw_step(np_inputs, np_attrs))
            store[f"{name}/step{t}/Grad"] = grad
            store[f"{name}/step{t}/ParamOut"] = np.asarray(state[name][0])
    store["hyper"] = np.array([learning_rate, weight_decay, beta1, beta2, 1e-8])
    np.savez_compressed(f"{OUT}/adamw_layerwise.npz", **store)


def _class_literals(path, cls_name, fn_name):
    """{attribute: literal} for the ``self.x = <literal or np.array([literal])...>`` lines of cls.fn."""
    tree = ast.parse(open(path).read())
    out = {}
    for cls in tree.body:
        if isinstance(cls, ast.ClassDef) and cls.name == cls_name:
            for fn in cls.body:
                if isinstance(fn, ast.FunctionDef) and fn.name == fn_name:
                    for node in ast.walk(fn):
                        if isinstance(node, ast.Assign) and isinstance(node.targets[0], ast.Attribute):
                            for sub in ast.walk(node.value):
                                if isinstance(sub, ast.Constant) and isinstance(sub.value, (int, float, bool)):
                                    out[node.targets[0].attr] = sub.value
                                    break
                        if isinstance(node, ast.Dict):
                            for k, v in zip(node.keys, node.values):
                                if isinstance(k, ast.Constant) and isinstance(v, ast.Constant):
                                    out[k.value] = v.value
    return out


def gen_update_loss_scaling():
    """Known answers of test_update_loss_scaling_op.py: TestUpdateLossScalingOp (:52-104, found_inf False at
    good = 999 of 1000 -> scale * incr_ratio, counters 0) and TestUpdateLossScalingOpBad (:107-136, found_inf True
    at bad = 1 of 2 -> scale * decr_ratio, counters 0, Out zeroed).  The literals are read from the AST of init()."""
    lit = _class_literals(f"{REF}/test_update_loss_scaling_op.py", "TestUpdateLossScalingOp", "init")
    need = ("incr_ratio", "decr_ratio", "prev_loss_scaling", "num_good_steps", "num_bad_steps",
            "incr_every_n_steps", "decr_every_n_nan_or_inf")
    assert all(k in lit for k in need), lit
    prev = np.array([lit["prev_loss_scaling"]]).astype(np.float32)
    cases = []
    for name, found, ratio, zero_out in (("TestUpdateLossScalingOp", False, lit["incr_ratio"], False),
                                         ("TestUpdateLossScalingOpBad", True, lit["decr_ratio"], True)):
        cases.append({
            "name": name, "found_inf": found, "prev_loss_scaling": float(prev[0]),
            "in_good_steps": lit["num_good_steps"], "in_bad_steps": lit["num_bad_steps"],
            "incr_every_n_steps": lit["incr_every_n_steps"], "decr_every_n_nan_or_inf": lit["decr_every_n_nan_or_inf"],
            "incr_ratio": lit["incr_ratio"], "decr_ratio": lit["decr_ratio"],
            "expected_loss_scaling": float((prev * ratio)[0]), "expected_good_steps": 0, "expected_bad_steps": 0,
            "expected_out_zeroed": zero_out,
        })
    with open(f"{OUT}/update_loss_scaling_cases.json", "w") as f:
        json.dump(cases, f, indent=1)


def gen_sparse_adam():
    """TestSparseAdamOp of test_adam_op.py (:355-411): literal inputs (height 10, rows [0, 4, 7], 12 columns, everything
    5.0, gradient ones with two exceptions), both lazy modes, through the reference's own ``adam_step_sparse`` (:304-352).
    The literals are reproduced from ``setup`` as written there."""
    step = extract_function(f"{REF}/test_adam_op.py", "adam_step_sparse")
    beta1, beta2, epsilon = 0.78, 0.836, 1e-4
    height, rows, row_numel = 10, [0, 4, 7], 12
    dense = {
        "Param": np.full((height, row_numel), 5.0).astype("float32"),
        "Moment1": np.full((height, row_numel), 5.0).astype("float32"),
        "Moment2": np.full((height, row_numel), 5.0).astype("float32"),
        "Beta1Pow": np.array([beta1**10]).astype("float32"),
        "Beta2Pow": np.array([beta2**10]).astype("float32"),
        "LearningRate": np.full((1), 2.0).astype("float32"),
    }
    attrs = {"epsilon": epsilon, "beta1": beta1, "beta2": beta2, "min_row_size_to_use_multithread": 2}
    grad = np.ones((len(rows), row_numel)).astype("float32")
    grad[0, 0] = 2.0
    grad[2, 8] = 4.0
    store = {"rows": np.array(rows, np.int64), "height": np.array([height]), "grad": grad,
             "hyper": np.array([beta1, beta2, epsilon])}
    for k, v in dense.items():
        store["in/" + k] = v
    for lazy in (False, True):
        p, m1, m2 = step(dense, attrs, height, rows, row_numel, grad, lazy)
        tag = "lazy" if lazy else "full"
        store[tag + "/ParamOut"], store[tag + "/Moment1Out"], store[tag + "/Moment2Out"] = p, m1, m2
    np.savez_compressed(f"{OUT}/sparse_adam_optest.npz", **store)


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("reference tree not present; fixtures can only be regenerated in the build container")
    gen_adamw()
    gen_adam()
    gen_check_finite()
    gen_squared_l2_norm()
    gen_bucket_cases()
    gen_fused_adam()
    gen_adamw_layerwise()
    gen_update_loss_scaling()
    gen_sparse_adam()
    print("fixtures written to", OUT)
