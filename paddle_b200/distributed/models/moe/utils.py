"""MoE op wrappers (reference: python/paddle/distributed/models/moe/utils.py:20-230)."""
from paddle_b200 import ops


def _number_count(numbers, upper_range):
    return ops.number_count(numbers, upper_range)


def _assign_pos(x, cum_count):
    return ops.assign_pos(x, cum_count)


def _random_routing(topk_idx, topk_value, prob, topk=2):
    return ops.random_routing(topk_idx, topk_value, prob, topk)


def _limit_by_capacity(expert_count, capacity, n_worker):
    return ops.limit_by_capacity(expert_count, capacity, n_worker)


def _prune_gate_by_capacity(gate_idx, expert_count, n_expert, n_worker):
    return ops.prune_gate_by_capacity(gate_idx, expert_count, n_expert, n_worker)
