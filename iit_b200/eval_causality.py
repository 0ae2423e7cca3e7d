"""The causality sweep: interchange accuracy of every (LL hook point, HL node) on counterfactual inputs.

API mirror of `evaluate_model_on_ablations` (reference: eval_causality.py:14-101). Per hook point the correspondence is
rebuilt (get_alignment with that hook point), per batch and HL node a counterfactual batch is generated that differs
from the base batch in that node's quadrant, the pair is interchanged, and
    stat[hook][node] = mean over batches of  sum(1[argmax(ll) == hl] * changed) / (sum(changed) + 1e-10),
    changed = (counterfactual label != base label)                                  (eval_causality.py:74-85).
Here: counterfactuals come from a device-resident digit bank (tasks/mnist_pvr/counterfactual.py), the masked counts
from the IIA kernel (integers, no host sync inside the loops), the source run stops at the hook point, and the counts
of all ranks are summed when a data-parallel group is given. One host read at the end."""
from __future__ import annotations

import numpy as np
import torch as t

from iit_b200 import ops
from iit_b200.model_pairs.iit_model_pair import IITModelPair
from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
from iit_b200.utils.wrapper import get_hook_points


def evaluate_model_on_ablations(ll_model, task, test_batches, eval_args, bank, verbose=False, dp=None, hook_points=None,
                                share_base_prefix=True):
    """`test_batches`: list of base inputs `(x, y, intermediate_vars)` already on the device (this rank's share);
    `bank`: DigitBank. Returns {hook_point: {hl_node: float}} like the reference, plus the evaluated pair count under
    the key "_pairs" when eval_args.get("count_pairs")."""
    test_batches = list(test_batches)
    if not test_batches:
        raise ValueError("no batches to evaluate")
    input_shape = tuple(test_batches[0][0].shape)
    hook_points = list(hook_points) if hook_points is not None else get_hook_points(ll_model)
    stats_per_layer, layout, pairs = {}, [], 0
    counts = []
    was_training = ll_model.training
    ll_model.eval()
    try:
        with t.no_grad():
            for hook_point in hook_points:
                _, hl_model, corr = get_alignment(task, config={"hook_point": hook_point, "input_shape": input_shape},
                                                  ll_model=ll_model)
                pair = IITModelPair(hl_model, ll_model, corr=corr)
                pair.stop_source_run_at_hook = True
                # the base run is the same for every node up to the residual block that holds the hook point: compute that
                # prefix once per batch and let each node's base run resume there (ResNet18LL.resume_point)
                rp = ll_model.resume_point(hook_point) if (share_base_prefix and hasattr(ll_model, "resume_point")) else None
                prefixes = None
                if rp is not None:
                    prefixes = [ll_model.run_with_cache(b[0], names_filter=[rp[1]], stop_when_cached=True)[1][rp[1]]
                                for b in test_batches]
                for hl_node in corr.keys():
                    per_batch = t.zeros((len(test_batches), 2), dtype=t.int64, device=test_batches[0][0].device)
                    for bi, base_input in enumerate(test_batches):
                        ablated = bank.patch_batch_at_hl(base_input[0], base_input[2], hl_node, base_input[1])
                        pair.base_resume = (rp[0], prefixes[bi]) if rp is not None else None
                        hl_output, ll_output = pair.do_intervention(base_input, ablated, hl_node)
                        changed = ablated[1] != base_input[1]
                        ops.iia_count(ll_output, hl_output, changed=changed, counts=per_batch[bi])
                        pairs += int(base_input[0].shape[0])
                    layout.append((hook_point, hl_node))
                    counts.append(per_batch)
    finally:
        ll_model.train(was_training)
    allc = t.stack(counts)                                   # [entries, batches, 2] (hits among changed, changed)
    if dp is not None and getattr(dp, "world", 1) > 1:
        allc = dp.allgather_counts(allc)                     # batches of every rank, concatenated along dim 1
    host = allc.cpu().numpy()
    for (hook_point, hl_node), c in zip(layout, host):
        # float32 arithmetic as torch does it on the reference's float tensors
        ratios = c[:, 0].astype(np.float32) / (c[:, 1].astype(np.float32) + np.float32(1e-10))
        val = float(ratios.astype(np.float32).sum(dtype=np.float32) / np.float32(len(ratios)))
        assert 0 <= val <= 1, f"hookpoint_stats[{hl_node}]: {val}"
        stats_per_layer.setdefault(hook_point, {})[hl_node] = val
        if verbose:
            print(f"hook_point: {hook_point} {hl_node}: {val}")
    if eval_args.get("count_pairs"):
        stats_per_layer["_pairs"] = pairs
    return stats_per_layer
