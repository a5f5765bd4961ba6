"""Presolve: host-side restatement of the reference's four reduction
rules and their deterministic queues, which `GeneralForm::standardize` ALWAYS runs before the simplex sees a problem
(/root/reference/src/data/linear_program/general_form/mod.rs:307-366).  Exact rationals throughout.

    Index / presolve_step          general_form/presolve/mod.rs:29-121, :122-295
    Counters                       presolve/counters.rs:14-108
    Queues (3 stacks + 1 deque)    presolve/queues.rs:15-119
    Updates, into_changes          presolve/updates.rs:17-420
    rules                          presolve/rule/fixed_variable.rs:21, bound_constraint.rs:27, slack.rs:64,
                                   domain_propagation.rs:34
    driver loop + applying changes general_form/mod.rs:368-391, :405-470, :728-790

The order in which rules fire decides which rows and columns survive, hence the standard form the simplex gets and
its pivot sequence -- so the queues are restated literally: `substitution`, `bound` and `slack` are LIFO stacks
(`Vec::pop`), `activity` is a FIFO deque with a membership set, rows are walked in column order and columns in row
order, and a queued index is skipped when its counter has dropped to zero meanwhile.
"""
from fractions import Fraction as Fr
from collections import deque

LOWER, UPPER = 0, 1
MEANINGFUL, NOT_MEANINGFUL, NONE = "meaningful", "not meaningful", "none"


class PresolveInfeasible(Exception):
    pass


class PresolveUnbounded(Exception):
    pass


def _sign(c):
    return 1 if c > 0 else -1


def _times(direction, sign):
    """BoundDirection * NonZeroSign (elements.rs): a negative coefficient swaps the direction."""
    return direction if sign > 0 else 1 - direction


def is_empty_constraint_feasible(rhs, ctype):
    """presolve/mod.rs:297-321"""
    if ctype == "E":
        return rhs == 0
    if ctype == "L":
        return rhs >= 0
    if ctype == "G":
        return rhs <= 0
    return rhs >= 0 and rhs - ctype[1] <= 0


def optimize_independent_column(maximize, cost, lower, upper):
    """updates.rs:398-420: a column that appears in no constraint sits at the bound its cost prefers."""
    if (not maximize and cost > 0) or (maximize and cost < 0):
        if lower is None:
            raise PresolveUnbounded()
        return lower
    if upper is None:
        raise PresolveUnbounded()
    return upper


class Index:
    """presolve/mod.rs:29-61.  `gf` needs: cols (column-major, entries sorted by row), cost, lower, upper, types, b, maximize."""

    def __init__(self, gf):
        self.gf = gf
        n, m = len(gf.cols), len(gf.b)
        # ---- Counters::new (counters.rs:31-66)
        self.rows = [[] for _ in range(m)]
        for j, col in enumerate(gf.cols):
            for i, v in col:
                self.rows[i].append((j, v))
        self.c_constraint = [len(r) for r in self.rows]
        self.c_variable = [len(c) for c in gf.cols]
        self.c_activity = []
        for row in self.rows:
            lo = up = 0
            for j, v in row:
                l, u = (gf.lower[j], gf.upper[j]) if v > 0 else (gf.upper[j], gf.lower[j])
                lo += l is None
                up += u is None
            self.c_activity.append([lo, up])
        # ---- Updates::new (updates.rs:41-95)
        self.u_b, self.u_constraints, self.u_bounds, self.u_activity_bounds = {}, {}, {}, {}
        self.fixed_cost = Fr(0)
        self.removed_variables = []
        for j in range(n):
            if self.c_variable[j] == 0:
                lo, up = gf.lower[j], gf.upper[j]
                if gf.cost[j] == 0:
                    value = up if up is not None else (lo if lo is not None else Fr(0))      # get_feasible_value, mod.rs:1055-1059
                else:
                    value = optimize_independent_column(gf.maximize, gf.cost[j], lo, up)
                    self.fixed_cost += gf.cost[j] * value
                self.removed_variables.append((j, ("S", value)))
        self.constraints_marked_removed = []
        for i in range(m):
            if self.c_constraint[i] == 0:
                if not is_empty_constraint_feasible(gf.b[i], gf.types[i]):
                    raise PresolveInfeasible()
                self.constraints_marked_removed.append(i)
        # ---- Queues::new (queues.rs:58-101)
        self.q_bound = [i for i in range(m) if self.c_constraint[i] == 1]
        self.q_activity, self.q_activity_set = deque(), set()
        for i in range(m):
            if self.c_constraint[i] > 1:
                if self.c_activity[i][0] <= 1:
                    self._activity_insert(i, LOWER)
                if self.c_activity[i][1] <= 1:
                    self._activity_insert(i, UPPER)
        self.q_slack = [j for j in range(n) if self.c_variable[j] == 1 and gf.cost[j] == 0]
        self.q_substitution = [j for j in range(n) if self.c_variable[j] > 0 and gf.lower[j] is not None and gf.lower[j] == gf.upper[j]]
        self.activity_bounds = [[None, None] for _ in range(m)]

    # ---- queues / counters ---------------------------------------------------------------------------
    def _activity_insert(self, constraint, direction):
        if (constraint, direction) not in self.q_activity_set:
            self.q_activity_set.add((constraint, direction))
            self.q_activity.append((constraint, direction))

    def are_queues_empty(self):
        return not (self.q_activity or self.q_slack or self.q_bound or self.q_substitution)

    def constraint_active(self, i):
        return self.c_constraint[i] > 0

    def variable_active(self, j):
        return self.c_variable[j] > 0

    def iter_active_column(self, j):
        return [(i, v) for (i, v) in self.gf.cols[j] if self.c_constraint[i] > 0]

    def iter_active_row(self, i):
        return [(j, v) for (j, v) in self.rows[i] if self.c_variable[j] > 0]

    # ---- Updates (updates.rs:96-260) -------------------------------------------------------------------
    def b(self, i):
        return self.u_b.get(i, self.gf.b[i])

    def change_b(self, i, change):
        self.u_b[i] = self.u_b.get(i, self.gf.b[i]) + change

    def constraint_type(self, i):
        return self.u_constraints.get(i, self.gf.types[i])

    def _original_bound(self, j, direction):
        return self.gf.lower[j] if direction == LOWER else self.gf.upper[j]

    def variable_bound(self, j, direction):
        key = (j, direction)
        if key in self.u_activity_bounds:
            return self.u_activity_bounds[key]
        if key in self.u_bounds:
            return self.u_bounds[key]
        return self._original_bound(j, direction)

    def is_variable_fixed(self, j):
        lo, up = self.variable_bound(j, LOWER), self.variable_bound(j, UPPER)
        return lo if (lo is not None and up is not None and lo == up) else None

    def variable_feasible_value(self, j):
        lo, up = self.variable_bound(j, LOWER), self.variable_bound(j, UPPER)
        if lo is None and up is None:
            return Fr(0)
        if lo is None or up is None:
            return up if lo is None else lo
        return up if lo <= up else None

    @staticmethod
    def _compare_and_update(key, new, existing, bounds):
        """updates.rs:375-396: ("shift", difference) when `new` is tighter, else None"""
        if (new > existing) if key[1] == LOWER else (new < existing):
            bounds[key] = new
            return ("shift", new - existing)
        return None

    def update_bound(self, j, direction, new):
        """updates.rs:166-204"""
        key = (j, direction)
        if key in self.u_bounds:
            existing = self.u_bounds[key]
        elif key in self.u_activity_bounds:
            existing = self.u_activity_bounds.pop(key)
            self.u_bounds[key] = existing
        else:
            existing = self._original_bound(j, direction)
            if existing is None:
                self.u_bounds[key] = new
                return ("new",)
        return self._compare_and_update(key, new, existing, self.u_bounds)

    def update_activity_variable_bound(self, j, direction, new):
        """updates.rs:205-246"""
        key = (j, direction)
        if key in self.u_activity_bounds:
            return self._compare_and_update(key, new, self.u_activity_bounds[key], self.u_activity_bounds)
        if key in self.u_bounds:
            return self._compare_and_update(key, new, self.u_bounds[key], self.u_bounds)
        original = self._original_bound(j, direction)
        if original is None:
            self.u_activity_bounds[key] = new
            return ("new",)
        return self._compare_and_update(key, new, original, self.u_activity_bounds)

    # ---- mod.rs:122-295 -----------------------------------------------------------------------------
    def presolve_step(self):
        if self.q_substitution:
            self.presolve_fixed_variable(self.q_substitution.pop())
            return MEANINGFUL
        while self.q_bound:
            i = self.q_bound.pop()
            if self.constraint_active(i):
                self.presolve_bound_constraint(i)
                return MEANINGFUL
        while self.q_slack:
            j = self.q_slack.pop()
            if self.variable_active(j):
                self.presolve_slack(j)
                return MEANINGFUL
        while self.q_activity:
            key = self.q_activity.popleft()
            self.q_activity_set.discard(key)
            if self.constraint_active(key[0]):
                return self.presolve_domain_propagation(*key)
        return NOT_MEANINGFUL

    def after_bound_change(self, j, direction, change):
        if self.is_variable_fixed(j) is not None and self.variable_active(j):
            self.q_substitution.append(j)
        if change is not None:
            self.update_activity_bounds(j, direction, change)
        else:
            self.update_activity_counters(j, direction)

    def update_activity_bounds(self, j, direction, by_how_much):
        for row, coefficient in self.iter_active_column(j):
            if not self.constraint_active(row):
                continue
            side = _times(direction, _sign(coefficient))
            if self.activity_bounds[row][side] is not None:
                self.activity_bounds[row][side] += by_how_much * coefficient
                self._activity_insert(row, side)

    def update_activity_counters(self, j, direction):
        for constraint, coefficient in self.iter_active_column(j):
            side = _times(direction, _sign(coefficient))
            self.c_activity[constraint][side] -= 1
            if self.c_activity[constraint][side] <= 1:
                self._activity_insert(constraint, side)

    def remove_constraint_values(self, constraint):
        for j, _ in self.iter_active_row(constraint):
            self.c_constraint[constraint] -= 1
            self.c_variable[j] -= 1
            self.queue_variable_by_counter(j)

    def queue_variable_by_counter(self, j):
        count = self.c_variable[j]
        if count == 0:
            if self.gf.cost[j] == 0:
                value = ("S", self.variable_feasible_value(j))
            else:
                v = optimize_independent_column(self.gf.maximize, self.gf.cost[j], self.variable_bound(j, LOWER), self.variable_bound(j, UPPER))
                self.fixed_cost += self.gf.cost[j] * v
                value = ("S", v)
            self.removed_variables.append((j, value))
        elif count == 1 and self.gf.cost[j] == 0:
            self.q_slack.append(j)

    def queue_constraint_by_counter(self, i):
        count = self.c_constraint[i]
        if count == 0:
            if not is_empty_constraint_feasible(self.b(i), self.constraint_type(i)):
                raise PresolveInfeasible()
            self.constraints_marked_removed.append(i)
            return MEANINGFUL
        if count == 1:
            self.q_bound.append(i)
        return NONE

    # ---- rule/fixed_variable.rs:21-48 ---------------------------------------------------------------------
    def presolve_fixed_variable(self, j):
        value = self.is_variable_fixed(j)
        for i, coefficient in self.iter_active_column(j):
            self.change_b(i, -coefficient * value)
        self.fixed_cost += self.gf.cost[j] * value
        for i, _ in self.iter_active_column(j):
            self.c_variable[j] -= 1
            self.c_constraint[i] -= 1
            self.queue_constraint_by_counter(i)
        self.removed_variables.append((j, ("S", value)))

    # ---- rule/bound_constraint.rs:27-85 -------------------------------------------------------------------
    def presolve_bound_constraint(self, constraint):
        (j, coefficient), = self.iter_active_row(constraint)
        bound_value = self.b(constraint) / coefficient
        ctype, positive = self.constraint_type(constraint), coefficient > 0
        if ctype == "E":
            changes = [(LOWER, bound_value), (UPPER, bound_value)]
        elif isinstance(ctype, tuple):
            bound1 = (self.b(constraint) - ctype[1]) / coefficient
            changes = [(LOWER, bound1), (UPPER, bound_value)] if positive else [(LOWER, bound_value), (UPPER, bound1)]
        elif (ctype == "G") == positive:
            changes = [(LOWER, bound_value)]
        else:
            changes = [(UPPER, bound_value)]
        self.c_variable[j] -= 1
        self.c_constraint[constraint] -= 1
        self.constraints_marked_removed.append(constraint)
        for direction, value in changes:
            ch = self.update_bound(j, direction, value)
            if ch is not None:
                self.after_bound_change(j, direction, None if ch[0] == "new" else ch[1])
        if self.variable_feasible_value(j) is None:
            raise PresolveInfeasible()
        self.queue_variable_by_counter(j)

    # ---- rule/slack.rs:64-214 -----------------------------------------------------------------------------
    def compute_removed_variable_solution(self, constraint, variable, coefficient):
        constant = self.b(constraint) / coefficient
        coefficients = [(self.gf.active_to_orig[j], other / coefficient) for (j, other) in self.iter_active_row(constraint) if j != variable]
        return ("F", constant, coefficients)

    def presolve_slack(self, variable):
        (constraint, coefficient), = self.iter_active_column(variable)
        ctype = self.constraint_type(constraint)
        lower, upper = self.variable_bound(variable, LOWER), self.variable_bound(variable, UPPER)
        none = (lower is None, upper is None)
        positive = coefficient > 0
        kind = "R" if isinstance(ctype, tuple) else ctype
        has = (lower is not None, upper is not None)
        remove_both = (has == (False, False) or
                       (kind == "G" and has == (True, False) and positive) or (kind == "L" and has == (False, True) and positive) or
                       (kind == "L" and has == (True, False) and not positive) or (kind == "G" and has == (False, True) and not positive))
        if remove_both:
            solution = self.compute_removed_variable_solution(constraint, variable, coefficient)
            for other, _ in self.iter_active_row(constraint):
                self.c_constraint[constraint] -= 1
                self.c_variable[other] -= 1
                if other != variable:
                    self.queue_variable_by_counter(other)
            self.removed_variables.append((variable, solution))
            self.constraints_marked_removed.append(constraint)
            return
        if has == (True, True) and kind in ("E", "R"):
            base = ctype[1] if kind == "R" else Fr(0)
            if positive:
                new_type, bound = ("R", base + coefficient * (upper - lower)), lower
            else:
                new_type, bound = ("R", base + coefficient * (lower - upper)), upper
        elif has == (True, True):                       # Less / Greater with both bounds: the side the relation leaves open
            if kind == "L":
                new_type, bound = ("L", lower) if positive else ("L", upper)
            else:
                new_type, bound = ("G", upper) if positive else ("G", lower)
        elif has == (True, False):
            new_type, bound = ("L", lower) if positive else ("G", lower)
        else:
            new_type, bound = ("G", upper) if positive else ("L", upper)
        change = -coefficient * bound
        if kind in ("E", "R"):
            removed = self.compute_removed_variable_solution(constraint, variable, coefficient)
        else:
            removed = ("S", bound)
        self.c_variable[variable] -= 1
        self.removed_variables.append((variable, removed))
        # update_activity_queues_if_needed (slack.rs:169-190)
        if (none[0] and positive) or (none[1] and not positive):
            self.c_activity[constraint][0] -= 1
            if self.c_activity[constraint][0] <= 1:
                self._activity_insert(constraint, LOWER)
        if (none[1] and positive) or (none[0] and not positive):
            self.c_activity[constraint][1] -= 1
            if self.c_activity[constraint][1] <= 1:
                self._activity_insert(constraint, UPPER)
        self.c_constraint[constraint] -= 1
        self.queue_constraint_by_counter(constraint)
        self.change_b(constraint, change)
        self.u_constraints[constraint] = new_type

    # ---- rule/domain_propagation.rs:34-487 ------------------------------------------------------------------
    def presolve_domain_propagation(self, constraint, direction):
        counter = self.c_activity[constraint][direction]
        if counter == 0:
            return self.for_entire_constraint(constraint, direction)
        if counter == 1:
            return self.create_variable_bound(constraint, direction)
        raise AssertionError("constraint queued before its activity counter reached 0 or 1")

    def compute_activity_bound_if_needed(self, constraint, direction):
        if self.activity_bounds[constraint][direction] is None:
            total = Fr(0)
            for j, coefficient in self.iter_active_row(constraint):
                total += coefficient * self.variable_bound(j, _times(direction, _sign(coefficient)))
            self.activity_bounds[constraint][direction] = total
        return self.activity_bounds[constraint][direction]

    def can_variable_rule_be_applied(self, constraint, direction):
        rhs, ctype = self.b(constraint), self.constraint_type(constraint)
        if ctype == "E":
            return rhs
        if isinstance(ctype, tuple):
            return rhs if direction == LOWER else rhs - ctype[1]
        if ctype == "L":
            return rhs if direction == LOWER else None
        return None if direction == LOWER else rhs

    def constraint_update(self, constraint, bound_value, direction):
        """-> None | "remove" | ("replace", relation, rhs shift) | "set" ; raises PresolveInfeasible (domain_propagation.rs:180-254)"""
        rhs, ctype = self.b(constraint), self.constraint_type(constraint)
        kind = "R" if isinstance(ctype, tuple) else ctype
        cmp = (rhs > bound_value) - (rhs < bound_value)
        if direction == LOWER:
            if kind in ("E", "R", "L") and cmp < 0:
                raise PresolveInfeasible()
            if kind in ("E", "L") and cmp == 0:
                return "set"
            if kind == "G" and cmp <= 0:
                return "remove"
            if kind == "R" and cmp > 0:
                return ("replace", "L", Fr(0)) if bound_value >= rhs - ctype[1] else None
            if kind == "R":
                raise AssertionError("a range of zero should be an equality")
            return None
        if kind in ("E", "G") and cmp > 0:
            raise PresolveInfeasible()
        if kind in ("E", "G") and cmp == 0:
            return "set"
        if kind == "L" and cmp >= 0:
            return "remove"
        if kind == "R" and cmp == 0:
            return ("replace", "G", -ctype[1])
        if kind == "R" and cmp > 0:
            lower_bound = rhs - ctype[1]
            if bound_value < lower_bound:
                raise PresolveInfeasible()
            return "set" if bound_value == lower_bound else ("replace", "G", -ctype[1])
        return None

    def for_entire_constraint(self, constraint, direction):
        change = [NONE]
        activity_bound = self.compute_activity_bound_if_needed(constraint, direction)
        update = self.constraint_update(constraint, activity_bound, direction)
        remove_constraint, apply_variable_part = False, True
        if update is not None:
            if update == "remove":
                remove_constraint = True
            elif update == "set":
                to_update = []
                for j, coefficient in self.iter_active_row(constraint):
                    vdir = _times(direction, _sign(coefficient))
                    bound = self.variable_bound(j, vdir)
                    if (j, vdir) in self.u_activity_bounds:
                        self.u_bounds[(j, vdir)] = self.u_activity_bounds.pop((j, vdir))
                    ch = self.update_bound(j, 1 - vdir, bound)
                    if ch is not None and ch[0] == "new":
                        to_update.append((j, 1 - vdir))
                    self.q_substitution.append(j)
                for j, d in to_update:
                    self.update_activity_counters(j, d)
                remove_constraint, apply_variable_part = True, False
            else:
                _, relation, shift = update
                self.u_constraints[constraint] = relation
                self.change_b(constraint, shift)
            change[0] = MEANINGFUL
        if apply_variable_part:
            rhs = self.can_variable_rule_be_applied(constraint, direction)
            if rhs is not None:
                self.variable_part(constraint, rhs, activity_bound, direction, change)
        if remove_constraint:
            self.remove_constraint_values(constraint)
            self.constraints_marked_removed.append(constraint)
        return change[0]

    def variable_part(self, constraint, rhs, activity_bound, activity_direction, change):
        for j, coefficient in self.iter_active_row(constraint):
            sign = _sign(coefficient)
            new_direction = _times(1 - activity_direction, sign)
            bound_value = self.variable_bound(j, _times(activity_direction, sign))
            residual = activity_bound - coefficient * bound_value
            new_value = (rhs - residual) / coefficient
            ch = self.update_activity_variable_bound(j, new_direction, new_value)
            if ch is None:
                continue
            if ch[0] == "new":
                self.after_bound_change(j, new_direction, None)
                change[0] = MEANINGFUL
            else:
                self.after_bound_change(j, new_direction, ch[1])
                if change[0] != MEANINGFUL:
                    change[0] = NOT_MEANINGFUL

    def create_variable_bound(self, constraint, activity_direction):
        rhs = self.can_variable_rule_be_applied(constraint, activity_direction)
        if rhs is None:
            return NONE
        total, target = Fr(0), None
        for j, coefficient in self.iter_active_row(constraint):
            bound = self.variable_bound(j, _times(activity_direction, _sign(coefficient)))
            if bound is None:
                if target is None:
                    target = (j, coefficient)
            else:
                total += coefficient * bound
        j, coefficient = target
        value = (rhs - total) / coefficient
        direction = _times(1 - activity_direction, _sign(coefficient))
        ch = self.update_activity_variable_bound(j, direction, value)
        if ch is None:
            return NONE
        if ch[0] == "new":
            self.after_bound_change(j, direction, None)
            return MEANINGFUL
        self.after_bound_change(j, direction, ch[1])
        return NOT_MEANINGFUL

    # ---- updates.rs:262-330 ----------------------------------------------------------------------------
    def into_changes(self):
        gf = self.gf
        removed_c = set(self.constraints_marked_removed)
        removed_v = {j for j, _ in self.removed_variables}
        b = {i: v for i, v in self.u_b.items() if i not in removed_c}
        constraints = {i: t for i, t in self.u_constraints.items() if i not in removed_c}
        bounds = {k: v for k, v in self.u_bounds.items() if k[0] not in removed_v}
        activity = {k: v for k, v in self.u_activity_bounds.items() if k[0] not in removed_v}
        # derived bounds are kept only when they turn an initially free variable into a bounded one
        free_to_be_restricted = {j for (j, _) in activity
                                 if gf.lower[j] is None and gf.upper[j] is None and (j, LOWER) not in bounds and (j, UPPER) not in bounds}
        for (j, d), v in activity.items():
            if j in free_to_be_restricted:
                bounds[(j, d)] = v
        b = {i: v for i, v in b.items() if v != gf.b[i]}
        constraints = {i: t for i, t in constraints.items() if t != gf.types[i]}
        return {"b": b, "constraints": constraints, "fixed_cost": self.fixed_cost, "bounds": bounds,
                "removed_variables": sorted(self.removed_variables, key=lambda x: x[0]),
                "constraints_marked_removed": sorted(self.constraints_marked_removed)}


def compute_presolve_changes(gf):
    """general_form/mod.rs:368-391"""
    index = Index(gf)
    without_meaningful = 0
    while not index.are_queues_empty() and without_meaningful < (len(gf.cols) - len(index.removed_variables)) + (len(gf.b) - len(index.constraints_marked_removed)):
        change = index.presolve_step()
        if change == MEANINGFUL:
            without_meaningful = 0
        elif change == NOT_MEANINGFUL:
            without_meaningful += 1
    return index.into_changes()
