def log_infeasible_constraints(*a, **k):
    pass


def log_infeasible_bounds(*a, **k):
    pass
