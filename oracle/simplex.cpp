// ORACLE (test infrastructure, not product) -- CPU restatement of RELP's two-phase revised
// simplex over Carry<RationalBig, BasisInverseRows<RationalBig>>.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  The product (rust-lp_b200/) never links, imports or calls it.
//
// The reference itself (Rust, mid-2021 nightly, crates not vendored) cannot be built in this
// image -- SURVEY.md section 0 F2-F4 -- so this single-threaded restatement is both the parity
// oracle and the timed CPU baseline ("kind": "port").  It keeps the reference's data structures
// (sparse sorted rows of B^-1 that are re-merged on every update, dense b and -pi, per-column
// pricing) so that its running time is representative of the algorithm the reference runs.
//
// Reference files followed (under /root/reference/src/algorithm/two_phase/):
//   tableau/kind/artificial/partially.rs:125-205      Partially::new
//   tableau/inverse_maintenance/carry/mod.rs:283-313  update_b
//                                           :326-333  update_minus_pi_and_obj
//                                           :381-426  create_for_partially_artificial
//                                           :214-271  create_minus_pi/obj_from_artificial
//                                           :484-510  from_artificial
//                                           :549-577  change_basis, cost_difference
//                                           :616-625  current_bfs
//                                           :650-689  from_artificial_remove_rows
//   .../carry/basis_inverse_rows.rs:42-83, 131-173, 194-205
//   data/linear_algebra/vector/sparse.rs:82-106, 213-248, 291-301 ; dense.rs:81-92
//   strategy/pivot_rule.rs:38-127 ; tableau/mod.rs:47-60, 102-126, 221-247
//   phase_one.rs:125-170, 223-260 ; phase_two.rs:22-51 ; mod.rs:30-75
//   matrix_provider/filter/generic_wrapper.rs:224-300 (rows removed from every column)
//
// Parity pins: src/tests/problem_1.rs, src/tests/problem_2.rs (full Carry states), the KATs in
// tableau/mod.rs:417-481 and pivot_rule.rs:137-167, tests/burkardt/test.rs exact objectives;
// all replayed in tests/test_oracle_*.py, and cross-checked pivot-for-pivot against the
// independent Python restatement oracle/pyref.py.
#include "rational.hpp"
#include <cstring>
#include <cstdio>
#include <chrono>

using namespace oracle;

typedef std::vector<std::pair<int, Rat>> SparseVec;   // ascending index, no zeros

struct LP {
    int m = 0, n = 0;
    std::vector<SparseVec> cols;   // all provider columns, virtual slacks materialised
    std::vector<Rat> cost;         // None == 0 (big/special.rs:50-60)
    std::vector<Rat> rhs;
    std::vector<std::pair<int, int>> pivots;   // pivot_element_indices(), ascending row
};

struct LogEntry { char phase; int j, r, leaving; Rat rc; int T, S; };

struct Carry {
    Rat minus_obj;
    std::vector<Rat> minus_pi, b;
    std::vector<int> basis_indices;
    std::vector<SparseVec> rows;
    int m() const { return (int)b.size(); }
};

// sparse.rs:82-106 (two-pointer merge of a sorted row with a sorted column)
static Rat sparse_inner_product(const SparseVec& row, const SparseVec& col) {
    Rat total;
    size_t i = 0;
    for (const auto& e : col) {
        while (i < row.size() && row[i].first < e.first) i++;
        if (i < row.size() && row[i].first == e.first) { total += row[i].second * e.second; i++; }
        if (i == row.size()) break;
    }
    return total;
}

// sparse.rs:213-248
static void add_multiple_of_row(SparseVec& self, const Rat& multiple, const SparseVec& other) {
    SparseVec out;
    out.reserve(self.size() + other.size());
    size_t j = 0;
    for (auto& e : self) {
        while (j < other.size() && other[j].first < e.first) {
            out.emplace_back(other[j].first, multiple * other[j].second); j++;
        }
        if (j < other.size() && other[j].first == e.first) {
            Rat nv = e.second + multiple * other[j].second;
            if (!nv.is_zero()) out.emplace_back(e.first, nv);
            j++;
        } else out.emplace_back(e.first, e.second);
    }
    for (; j < other.size(); j++) out.emplace_back(other[j].first, multiple * other[j].second);
    self.swap(out);
}

static const Rat* sparse_get(const SparseVec& v, int idx) {
    size_t lo = 0, hi = v.size();
    while (lo < hi) { size_t mid = (lo + hi) / 2; if (v[mid].first < idx) lo = mid + 1; else hi = mid; }
    return (lo < v.size() && v[lo].first == idx) ? &v[lo].second : nullptr;
}

struct Solver {
    LP lp;            // current provider (rows possibly removed after phase 1)
    Carry c;
    std::vector<char> in_basis;
    bool phase1 = true;
    std::vector<int> column_to_row;
    int nr_art = 0;
    int rule1 = 1, rule2 = 2;
    int last_selected = -1;          // FirstProfitableWithMemory state (pivot_rule.rs:62-94)
    int stage = 0;                   // 0 = phase 1 loop, 1 = phase 2 loop, 2 = finished
    int status = 0;                  // 0 running, 1 optimal, 2 unbounded, 3 infeasible
    std::vector<LogEntry> log;
    std::vector<int> rows_removed;
    std::string scratch;
    double loop_seconds = 0;

    int nr_columns() const { return phase1 ? nr_art + lp.n : lp.n; }
    // partially.rs:72-80 / non_artificial.rs:58-62 -- cloned per call like matrix_data.rs:318
    SparseVec original_column(int j) const {
        if (phase1) {
            if (j < nr_art) return SparseVec{{column_to_row[j], Rat(1)}};
            return lp.cols[j - nr_art];
        }
        return lp.cols[j];
    }
    // tableau/mod.rs:102-108 with carry/mod.rs:572-577 and dense.rs:81-92
    Rat relative_cost(int j) const {
        SparseVec col = original_column(j);
        Rat total;
        for (const auto& e : col) total += c.minus_pi[e.first] * e.second;
        if (phase1) return j < nr_art ? total + Rat(1) : total;   // Cost::One / Cost::Zero
        return total + lp.cost[j];
    }
    // basis_inverse_rows.rs:144-173
    Rat generate_element(int i, const SparseVec& col) const { return sparse_inner_product(c.rows[i], col); }
    SparseVec generate_column(int j) const {
        SparseVec out;
        for (int i = 0; i < c.m(); i++) {
            SparseVec col = original_column(j);   // the reference clones a_j per row (:149)
            Rat e = generate_element(i, col);
            if (!e.is_zero()) out.emplace_back(i, e);
        }
        return out;
    }
    // tableau/mod.rs:221-247
    int select_primal_pivot_row(const SparseVec& col) const {
        int min_index = -1, min_leaving = 0; Rat min_ratio;
        for (const auto& e : col) {
            if (e.second.sign() > 0) {
                Rat ratio = c.b[e.first] / e.second;
                int leaving = c.basis_indices[e.first];
                if (min_index >= 0) {
                    if (ratio == min_ratio && leaving < min_leaving) { min_index = e.first; min_leaving = leaving; }
                    else if (ratio < min_ratio) { min_index = e.first; min_ratio = ratio; min_leaving = leaving; }
                } else { min_index = e.first; min_ratio = ratio; min_leaving = leaving; }
            }
        }
        return min_index;
    }
    // carry/mod.rs:549-570
    int change_basis(int r, int j, const SparseVec& col, const Rat& rc) {
        const Rat* pv = sparse_get(col, r);
        if (!pv) { fprintf(stderr, "Pivot value can't be zero.\n"); abort(); }
        Rat pivot = *pv;
        // update_b (283-313)
        c.b[r] = c.b[r] / pivot;
        for (const auto& e : col) if (e.first != r) c.b[e.first] -= e.second * c.b[r];
        // normalize_pivot_row (42-51) -> element_wise_divide (sparse.rs:291-301)
        for (auto& e : c.rows[r]) e.second = e.second / pivot;
        // row_reduce (63-83)
        for (const auto& e : col) if (e.first != r) add_multiple_of_row(c.rows[e.first], -e.second, c.rows[r]);
        // update_minus_pi_and_obj (326-333)
        SparseVec prow = c.rows[r];   // basis_inverse_row clones (:181-183)
        for (const auto& e : prow) c.minus_pi[e.first] -= rc * e.second;
        c.minus_obj -= rc * c.b[r];
        int leaving = c.basis_indices[r];
        c.basis_indices[r] = j;
        return leaving;
    }
    // tableau/mod.rs:47-84
    void bring_into_basis(int j, int r, const SparseVec& col, const Rat& rc, char tag) {
        int T = (int)col.size() - 1, S = (int)c.rows[r].size() + 1;
        int leaving = change_basis(r, j, col, rc);
        in_basis[leaving] = 0; in_basis[j] = 1;
        log.push_back(LogEntry{tag, j, r, leaving, rc, T, S});
    }

    // pivot_rule.rs:38-127
    bool find_first(int lo, int hi, int& j_out, Rat& rc_out) const {
        for (int j = lo; j < hi; j++) {
            if (in_basis[j]) continue;
            Rat rc = relative_cost(j);
            if (rc.sign() < 0) { j_out = j; rc_out = rc; return true; }
        }
        return false;
    }
    bool select_column(int rule, int& j_out, Rat& rc_out) {
        int n = nr_columns();
        if (rule == 0) return find_first(0, n, j_out, rc_out);
        if (rule == 1) {
            bool ok;
            if (last_selected < 0) ok = find_first(0, n, j_out, rc_out);
            else ok = find_first(last_selected, n, j_out, rc_out) || find_first(0, last_selected, j_out, rc_out);
            last_selected = ok ? j_out : -1;
            return ok;
        }
        bool have = false;   // SteepestDescent: strict <, ties keep the lowest j
        for (int j = 0; j < n; j++) {
            if (in_basis[j]) continue;
            Rat rc = relative_cost(j);
            if (rc.sign() < 0 && (!have || rc < rc_out)) { have = true; j_out = j; rc_out = rc; }
        }
        return have;
    }

    // partially.rs:125-205 + carry/mod.rs:381-426
    void init_partially_artificial() {
        int m = lp.m;
        std::vector<int> real_col(m, -1);
        for (auto& p : lp.pivots) real_col[p.first] = p.second;
        column_to_row.clear();
        for (int i = 0; i < m; i++) if (real_col[i] < 0) column_to_row.push_back(i);
        nr_art = (int)column_to_row.size();
        c.basis_indices.assign(m, 0);
        c.minus_pi.assign(m, Rat(0));
        c.minus_obj = Rat(0);
        int k = 0;
        for (int i = 0; i < m; i++) {
            if (real_col[i] < 0) { c.basis_indices[i] = k++; c.minus_pi[i] = Rat(-1); c.minus_obj -= lp.rhs[i]; }
            else c.basis_indices[i] = nr_art + real_col[i];
        }
        c.b = lp.rhs;
        c.rows.assign(m, SparseVec());
        for (int i = 0; i < m; i++) c.rows[i].emplace_back(i, Rat(1));
        phase1 = true;
        in_basis.assign(nr_art + lp.n, 0);
        for (int i = 0; i < m; i++) in_basis[c.basis_indices[i]] = 1;
        last_selected = -1; stage = 0; status = 0; log.clear(); rows_removed.clear();
    }

    // phase_one.rs:223-260
    bool cleanup_any = false;      // test/diagnostic switch: pivot on any nonzero element of the row (see relp_set_option "cleanup_any")
    void remove_artificial_basis_variables() {
        rows_removed.clear();
        std::vector<int> arts;   // snapshot, ascending (phase_one.rs:230-232)
        for (int a = 0; a < nr_art; a++) if (in_basis[a]) arts.push_back(a);
        for (int a : arts) {
            int pivot_row = column_to_row[a];       // the reference: the row the artificial STARTED on (partially.rs:98-102)
            if (cleanup_any)                        // the row that holds it NOW (it may have left and re-entered on another row)
                for (int i = 0; i < c.m(); i++) if (c.basis_indices[i] == a) pivot_row = i;
            int found = -1; Rat cost;
            for (int j = nr_art; j < nr_columns(); j++) {
                if (in_basis[j]) continue;
                Rat rc = relative_cost(j);
                if (!cleanup_any && !rc.is_zero()) continue;       // the reference tries zero-cost columns only (phase_one.rs:243)
                if (!generate_element(pivot_row, original_column(j)).is_zero()) { found = j; cost = rc; break; }
            }
            if (found >= 0) {
                SparseVec col = generate_column(found);
                bring_into_basis(found, pivot_row, col, cost, 'z');
            } else if (pivot_row == column_to_row[a]) rows_removed.push_back(pivot_row);
            // (cleanup_any, artificial on a foreign row whose real part is all zero: it stays, the hand-over reports it)
            // the reference pushes `artificial` (phase_one.rs:251); equal to pivot_row whenever the reference survives (else it removes the wrong row and underflows at carry/mod.rs:665)
        }
    }

    // carry/mod.rs:214-271 (the reference regenerates B^-1 column by column; same values)
    void recompute_cost_row() {
        int m = c.m();
        std::vector<Rat> pi(m);
        for (int i = 0; i < m; i++) {
            const Rat& cb = lp.cost[c.basis_indices[i]];
            for (const auto& e : c.rows[i]) pi[e.first] += e.second * cb;
        }
        for (int k = 0; k < m; k++) c.minus_pi[k] = -pi[k];
        Rat obj;
        for (int i = 0; i < m; i++) obj += c.b[i] * lp.cost[c.basis_indices[i]];
        c.minus_obj = -obj;
    }

    // non_artificial.rs:151-220 ; carry/mod.rs:484-510, 650-689 ; generic_wrapper.rs:224-300
    void to_phase_two() {
        std::vector<char> basis = in_basis;
        if (!rows_removed.empty()) {
            int m = c.m();
            for (int row : rows_removed) basis[c.basis_indices[row]] = 0;
            std::vector<int> newidx(m, -1); int k = 0; size_t q = 0;
            for (int i = 0; i < m; i++) {
                if (q < rows_removed.size() && rows_removed[q] == i) { q++; continue; }
                newidx[i] = k++;
            }
            Carry nc; nc.basis_indices.resize(k); nc.b.resize(k); nc.rows.resize(k); nc.minus_pi.resize(k);
            for (int i = 0; i < m; i++) {
                if (newidx[i] < 0) continue;
                nc.basis_indices[newidx[i]] = c.basis_indices[i];
                nc.b[newidx[i]] = c.b[i];
                for (auto& e : c.rows[i]) if (newidx[e.first] >= 0) nc.rows[newidx[i]].emplace_back(newidx[e.first], e.second);
            }
            c = nc;
            LP nl; nl.m = k; nl.n = lp.n; nl.cost = lp.cost; nl.cols.resize(lp.n); nl.rhs.resize(k);
            for (int i = 0; i < m; i++) if (newidx[i] >= 0) nl.rhs[newidx[i]] = lp.rhs[i];
            for (int j = 0; j < lp.n; j++)
                for (auto& e : lp.cols[j]) if (newidx[e.first] >= 0) nl.cols[j].emplace_back(newidx[e.first], e.second);
            lp = nl;
        }
        // An artificial that re-entered the basis during phase 1 on another row is not found by the clean-up
        // (it pivots on the artificial's ORIGINAL row, partially.rs pivot_row_from_artificial); the reference
        // then underflows `basis_column -= nr_artificial` (carry/mod.rs:500/:665).  Reported as status -8.
        for (int x : c.basis_indices) if (x < nr_art) { status = -8; return; }
        for (auto& x : c.basis_indices) x -= nr_art;
        phase1 = false;
        recompute_cost_row();
        in_basis.assign(lp.n, 0);
        for (int j = nr_art; j < (int)basis.size(); j++) if (basis[j]) in_basis[j - nr_art] = 1;
        last_selected = -1;
    }

    // carry/mod.rs:428-463 (from_basis): basis[i] = column in row i.  The reference inverts through LU
    // (basis_inverse_rows.rs:103-129); the inverse is unique, here plain Gauss-Jordan on [B | I].
    bool from_basis(const std::vector<int>& basis) {
        int m = lp.m;
        std::vector<std::vector<Rat>> M(m, std::vector<Rat>(2 * m));
        for (int k = 0; k < m; k++) {
            for (const auto& e : lp.cols[basis[k]]) M[e.first][k] = e.second;
            M[k][m + k] = Rat(1);
        }
        for (int k = 0; k < m; k++) {
            int p = -1;
            for (int i = k; i < m; i++) if (!M[i][k].is_zero()) { p = i; break; }
            if (p < 0) return false;
            std::swap(M[p], M[k]);
            Rat piv = M[k][k];
            for (int q = 0; q < 2 * m; q++) if (!M[k][q].is_zero()) M[k][q] = M[k][q] / piv;
            for (int i = 0; i < m; i++) {
                if (i == k || M[i][k].is_zero()) continue;
                Rat f = M[i][k];
                for (int q = 0; q < 2 * m; q++) if (!M[k][q].is_zero()) M[i][q] -= f * M[k][q];
            }
        }
        nr_art = 0; phase1 = false; stage = 1; status = 0; last_selected = -1; log.clear(); rows_removed.clear();
        c.basis_indices = basis; c.rows.assign(m, SparseVec()); c.b.assign(m, Rat()); c.minus_pi.assign(m, Rat());
        for (int i = 0; i < m; i++) {
            Rat bi;
            for (int k = 0; k < m; k++) {
                const Rat& v = M[i][m + k];
                if (v.is_zero()) continue;
                c.rows[i].emplace_back(k, v);
                bi += v * lp.rhs[k];
            }
            c.b[i] = bi;
        }
        recompute_cost_row();
        in_basis.assign(lp.n, 0);
        for (int j : basis) in_basis[j] = 1;
        return true;
    }

    // phase_one.rs:125-170, phase_two.rs:22-51, mod.rs:30-75 as a resumable state machine
    int advance(long max_new_pivots, bool stop_at_handover = false) {
        auto t0 = std::chrono::steady_clock::now();
        long done = 0;
        while (status == 0) {
            if (max_new_pivots >= 0 && done >= max_new_pivots) break;
            int j; Rat rc;
            bool ok = select_column(stage == 0 ? rule1 : rule2, j, rc);
            if (!ok) {
                if (stage == 0) {
                    if (!c.minus_obj.is_zero()) { status = 3; break; }
                    bool art = false;
                    for (int a = 0; a < nr_art; a++) if (in_basis[a]) art = true;
                    if (art) remove_artificial_basis_variables();
                    to_phase_two();
                    if (status != 0) break;
                    stage = 1;
                    if (stop_at_handover) break;
                    continue;
                }
                status = 1; break;
            }
            SparseVec col = generate_column(j);
            int r = select_primal_pivot_row(col);
            if (r < 0) {
                if (stage == 0) { fprintf(stderr, "Artificial cost can not be unbounded.\n"); abort(); }
                status = 2; break;
            }
            bring_into_basis(j, r, col, rc, stage == 0 ? '1' : '2');
            done++;
        }
        loop_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        return status;
    }
};

// ---------------------------------------------------------------------------------------------
// C interface for ctypes (tests/, bench.py cpu_baseline).  Rationals go in as int64 num/den and
// come out as decimal strings "num/den" (or "num"); strings stay valid until the next call.
// ---------------------------------------------------------------------------------------------
extern "C" {

int orc_from_basis(void* h, const int32_t* basis);

void* orc_create(int m, int n, const int64_t* col_ptr, const int32_t* row_idx,
                 const int64_t* val_num, const int64_t* val_den,
                 const int64_t* cost_num, const int64_t* cost_den,
                 const int64_t* rhs_num, const int64_t* rhs_den,
                 int n_piv, const int32_t* piv_rows, const int32_t* piv_cols) {
    Solver* s = new Solver();
    LP& lp = s->lp;
    lp.m = m; lp.n = n; lp.cols.resize(n); lp.cost.resize(n); lp.rhs.resize(m);
    for (int j = 0; j < n; j++) {
        for (int64_t k = col_ptr[j]; k < col_ptr[j + 1]; k++)
            lp.cols[j].emplace_back(row_idx[k], Rat(BigInt(val_num[k]), BigInt(val_den[k])));
        lp.cost[j] = Rat(BigInt(cost_num[j]), BigInt(cost_den[j]));
    }
    for (int i = 0; i < m; i++) lp.rhs[i] = Rat(BigInt(rhs_num[i]), BigInt(rhs_den[i]));
    for (int k = 0; k < n_piv; k++) lp.pivots.emplace_back(piv_rows[k], piv_cols[k]);
    return s;
}
void orc_destroy(void* h) { delete (Solver*)h; }
void orc_set_cleanup_any(void* h, int v) { ((Solver*)h)->cleanup_any = v != 0; }

// Overwrite one matrix value / cost / rhs with an arbitrary-precision rational (decimal string).
void orc_set_value(void* h, int j, int k, const char* v) { ((Solver*)h)->lp.cols[j][k].second = Rat::from_string(v); }
void orc_set_cost(void* h, int j, const char* v) { ((Solver*)h)->lp.cost[j] = Rat::from_string(v); }
void orc_set_rhs(void* h, int i, const char* v) { ((Solver*)h)->lp.rhs[i] = Rat::from_string(v); }

void orc_begin(void* h, int rule1, int rule2) {
    Solver* s = (Solver*)h; s->rule1 = rule1; s->rule2 = rule2; s->loop_seconds = 0;
    s->init_partially_artificial();
}
// returns 0 running (pivot budget reached), 1 optimal, 2 unbounded, 3 infeasible
int orc_advance(void* h, long max_new_pivots) { return ((Solver*)h)->advance(max_new_pivots); }
// run phase 1 (and the artificial clean-up + hand-over) only; no phase-2 pivot is taken
int orc_advance_to_handover(void* h) { Solver* s = (Solver*)h; return s->stage == 0 ? s->advance(-1, true) : s->status; }
int orc_status(void* h) { return ((Solver*)h)->status; }
int orc_stage(void* h) { return ((Solver*)h)->stage; }
int orc_m(void* h) { return ((Solver*)h)->c.m(); }
int orc_nr_artificial(void* h) { return ((Solver*)h)->nr_art; }
long orc_nr_pivots(void* h) { return (long)((Solver*)h)->log.size(); }
double orc_loop_seconds(void* h) { return ((Solver*)h)->loop_seconds; }
int orc_nr_rows_removed(void* h) { return (int)((Solver*)h)->rows_removed.size(); }

// log entry k: phase tag ('1','2','z'), j, r, leaving, T, S ; rc via orc_log_rc
void orc_log_entry(void* h, long k, int32_t* out6) {
    const LogEntry& e = ((Solver*)h)->log[k];
    out6[0] = e.phase; out6[1] = e.j; out6[2] = e.r; out6[3] = e.leaving; out6[4] = e.T; out6[5] = e.S;
}
void orc_log_bulk(void* h, int32_t* out) {   // 6 int32 per entry
    Solver* s = (Solver*)h;
    for (size_t k = 0; k < s->log.size(); k++) orc_log_entry(h, (long)k, out + 6 * k);
}
const char* orc_log_rc(void* h, long k) { Solver* s = (Solver*)h; s->scratch = s->log[k].rc.to_string(); return s->scratch.c_str(); }

const char* orc_objective(void* h) { Solver* s = (Solver*)h; s->scratch = (-s->c.minus_obj).to_string(); return s->scratch.c_str(); }
void orc_basis_indices(void* h, int32_t* out) { Solver* s = (Solver*)h; for (int i = 0; i < s->c.m(); i++) out[i] = s->c.basis_indices[i]; }
const char* orc_b(void* h, int i) { Solver* s = (Solver*)h; s->scratch = s->c.b[i].to_string(); return s->scratch.c_str(); }
const char* orc_minus_pi(void* h, int i) { Solver* s = (Solver*)h; s->scratch = s->c.minus_pi[i].to_string(); return s->scratch.c_str(); }
// Whole Carry as text: line 0 "-obj", line 1 "-pi ...", line 2 "b ...", line 3 basis, then one
// line per B^-1 row "k:v k:v ..." (sparse, ascending).
const char* orc_dump_carry(void* h) {
    Solver* s = (Solver*)h; std::string& o = s->scratch; o.clear();
    o += s->c.minus_obj.to_string(); o += "\n";
    for (auto& v : s->c.minus_pi) { o += v.to_string(); o += " "; } o += "\n";
    for (auto& v : s->c.b) { o += v.to_string(); o += " "; } o += "\n";
    for (int v : s->c.basis_indices) { o += std::to_string(v); o += " "; } o += "\n";
    for (auto& row : s->c.rows) {
        for (auto& e : row) { o += std::to_string(e.first); o += ":"; o += e.second.to_string(); o += " "; }
        o += "\n";
    }
    return o.c_str();
}
// max bit length over all stored numerators/denominators (escalation statistics)
long orc_max_bits(void* h) {
    Solver* s = (Solver*)h; size_t mx = 0;
    auto upd = [&](const Rat& r) { mx = std::max(mx, std::max(r.n.bits(), r.d.bits())); };
    upd(s->c.minus_obj);
    for (auto& v : s->c.minus_pi) upd(v);
    for (auto& v : s->c.b) upd(v);
    for (auto& row : s->c.rows) for (auto& e : row) upd(e.second);
    return (long)mx;
}
long orc_nnz_inverse(void* h) { Solver* s = (Solver*)h; long t = 0; for (auto& r : s->c.rows) t += (long)r.size(); return t; }

// Rational KAT helper: op in "+-*/<" on decimal strings (tests/test_oracle_rational.py)
const char* orc_rat_op(const char* a, const char* b, char op) {
    static std::string out;
    Rat x = Rat::from_string(a), y = Rat::from_string(b);
    switch (op) {
        case '+': out = (x + y).to_string(); break;
        case '-': out = (x - y).to_string(); break;
        case '*': out = (x * y).to_string(); break;
        case '/': out = (x / y).to_string(); break;
        case '<': out = std::to_string(Rat::cmp(x, y)); break;
        default: out = "?";
    }
    return out.c_str();
}
int orc_from_basis(void* h, const int32_t* basis) {
    Solver* s = (Solver*)h;
    return s->from_basis(std::vector<int>(basis, basis + s->lp.m)) ? 0 : -1;
}

}
