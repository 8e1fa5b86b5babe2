// quotient.cuh -- the pointwise part of Evaluator::evaluate_h (halo2_backend/src/plonk/evaluation.rs:317-623): after the
// coset NTTs, every row of the extended domain is folded independently with y.  
//
//   graph_evaluate     GraphEvaluator::evaluate (evaluation.rs:789-834) over all rows: a small interpreter for the
//                      reference's `Calculation` / `ValueSource` program (evaluation.rs:26-172); one thread per row,
//                      intermediates in a [n_intermediates][threads] scratch matrix (coalesced), rotations resolved once per row.
//   permutation_fold   evaluation.rs:408-493     lookup_fold   evaluation.rs:517-565     shuffle_fold   evaluation.rs:573-620
//
// Inputs are the device-resident extended-domain columns produced by coeff_to_extended; nothing crosses PCIe.
// Per row the work is a few dozen field multiplications against a few dozen 32-byte loads: integer-pipe bound like the NTT.
#pragma once
#include "field.cuh"
#include "field_wide.cuh"

namespace h2b {

// ---- program encoding (what the Rust shim serialises from GraphEvaluator; evaluation.rs:26-47, 101-121) ----------------
// A value source is two words: w0 = kind | (rotation_index << 8), w1 = index (constant / intermediate / column / challenge).
enum : uint32_t {
    QS_CONSTANT = 0, QS_INTERMEDIATE = 1, QS_FIXED = 2, QS_ADVICE = 3, QS_INSTANCE = 4, QS_CHALLENGE = 5,
    QS_BETA = 6, QS_GAMMA = 7, QS_THETA = 8, QS_Y = 9, QS_PREVIOUS = 10
};
// A calculation is [op, target, a.w0, a.w1, b.w0, b.w1]; Horner is [op, target, start.w0, start.w1, factor.w0, factor.w1,
// n_parts, part0.w0, part0.w1, ...].
enum : uint32_t { QC_ADD = 0, QC_SUB = 1, QC_MUL = 2, QC_SQUARE = 3, QC_DOUBLE = 4, QC_NEGATE = 5, QC_HORNER = 6, QC_STORE = 7 };
constexpr uint32_t Q_MAX_ROTATIONS = 32;

struct GraphArgs {
    const uint32_t* prog;         // device
    uint32_t n_words, n_rot, n_inter, last_target;
    int32_t has_calcs;
    const Fr* constants;          // device
    const int32_t* rotations;     // device
    const Fr* const* fixed;       // device arrays of device column pointers
    const Fr* const* advice;
    const Fr* const* instance;
    const Fr* challenges;         // device
    Fr beta, gamma, theta, y;
    const Fr* prev;               // previous value per row, or nullptr (zero)
    Fr* out;
    Fr* scratch;                  // n_inter * n_threads
    uint32_t size, rot_scale;
};

struct PermFoldArgs {
    Fr* values;                               // in/out, `size` rows
    const Fr* const* sets;                    // permutation_product_coset of every set
    const Fr* const* col_values;              // the permuted columns' cosets, in p.columns order
    const Fr* const* col_perms;               // pk.permutation.cosets, same order
    uint32_t n_sets, n_cols, chunk_len;
    const Fr *l0, *l_last, *l_active;
    Fr beta, gamma, y, delta_start, delta, ext_omega, one;
    int32_t last_rotation;
    uint32_t size, rot_scale;
};

struct LookupFoldArgs {
    Fr* values;
    const Fr *table_values, *product, *permuted_input, *permuted_table, *l0, *l_last, *l_active;
    Fr beta, gamma, y, one;
    uint32_t size, rot_scale;
};

struct ShuffleFoldArgs {
    Fr* values;
    const Fr *input_values, *shuffle_values, *product, *l0, *l_last, *l_active;
    Fr y, one;
    uint32_t size, rot_scale;
};

// evaluation.rs:21-24 get_rotation_idx: (idx + rot * rot_scale) mod isize, Euclidean
H2B_HD uint32_t rotation_idx(uint32_t idx, int32_t rot, uint32_t rot_scale, uint32_t size) {
    int64_t v = (int64_t)idx + (int64_t)rot * (int64_t)rot_scale;
    int64_t m = v % (int64_t)size;
    if (m < 0) m += size;
    return (uint32_t)m;
}

#if defined(__CUDACC__) && !defined(H2B_NO_KERNELS)

H2B_D Fr q_fetch(const GraphArgs& A, uint32_t w0, uint32_t w1, const uint32_t* rot_idx, uint32_t gtid, uint32_t nthreads,
                 const Fr& prev) {
    switch (w0 & 0xffu) {
        case QS_CONSTANT: return fe_load(A.constants + w1);
        case QS_INTERMEDIATE: return fe_load(A.scratch + (size_t)w1 * nthreads + gtid);
        case QS_FIXED: return fe_load(A.fixed[w1] + rot_idx[w0 >> 8]);
        case QS_ADVICE: return fe_load(A.advice[w1] + rot_idx[w0 >> 8]);
        case QS_INSTANCE: return fe_load(A.instance[w1] + rot_idx[w0 >> 8]);
        case QS_CHALLENGE: return fe_load(A.challenges + w1);
        case QS_BETA: return A.beta;
        case QS_GAMMA: return A.gamma;
        case QS_THETA: return A.theta;
        case QS_Y: return A.y;
        default: return prev;
    }
}

__global__ void __launch_bounds__(128)
quotient_graph_kernel(const GraphArgs A) {
    const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t nthreads = gridDim.x * blockDim.x;
    uint32_t rot_idx[Q_MAX_ROTATIONS];
    for (uint32_t idx = gtid; idx < A.size; idx += nthreads) {
        for (uint32_t r = 0; r < A.n_rot; r++) rot_idx[r] = rotation_idx(idx, A.rotations[r], A.rot_scale, A.size);
        Fr prev = A.prev ? fe_load(A.prev + idx) : Fr::zero();
        uint32_t pc = 0;
        while (pc < A.n_words) {
            const uint32_t op = A.prog[pc], target = A.prog[pc + 1];
            Fr a = q_fetch(A, A.prog[pc + 2], A.prog[pc + 3], rot_idx, gtid, nthreads, prev);
            Fr r;
            if (op == QC_HORNER) {
                Fr factor = q_fetch(A, A.prog[pc + 4], A.prog[pc + 5], rot_idx, gtid, nthreads, prev);
                const uint32_t np = A.prog[pc + 6];
                r = a;
                for (uint32_t i = 0; i < np; i++) {
                    Fr part = q_fetch(A, A.prog[pc + 7 + 2 * i], A.prog[pc + 8 + 2 * i], rot_idx, gtid, nthreads, prev);
                    r = fe_add(fe_mul(r, factor), part);
                }
                pc += 7 + 2 * np;
            } else {
                if (op <= QC_MUL) {
                    Fr b = q_fetch(A, A.prog[pc + 4], A.prog[pc + 5], rot_idx, gtid, nthreads, prev);
                    r = op == QC_ADD ? fe_add(a, b) : op == QC_SUB ? fe_sub(a, b) : fe_mul(a, b);
                } else if (op == QC_SQUARE) r = fe_sqr_w(a);
                else if (op == QC_DOUBLE) r = fe_dbl(a);
                else if (op == QC_NEGATE) r = fe_neg(a);
                else r = a;
                pc += 6;
            }
            fe_store(A.scratch + (size_t)target * nthreads + gtid, r);
        }
        Fr res = A.has_calcs ? fe_load(A.scratch + (size_t)A.last_target * nthreads + gtid) : Fr::zero();
        fe_store(A.out + idx, res);
    }
}

// evaluation.rs:408-493
__global__ void __launch_bounds__(128)
quotient_permutation_kernel(const PermFoldArgs A) {
    const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t nthreads = gridDim.x * blockDim.x;
    if (gtid >= A.size) return;
    Fr beta_term = fe_pow_u64(A.ext_omega, (uint64_t)gtid);                  // extended_omega^idx (:411)
    const Fr stride = fe_pow_u64(A.ext_omega, (uint64_t)nthreads);
    for (uint32_t idx = gtid; idx < A.size; idx += nthreads) {
        const uint32_t r_next = rotation_idx(idx, 1, A.rot_scale, A.size);
        const uint32_t r_last = rotation_idx(idx, A.last_rotation, A.rot_scale, A.size);
        Fr value = fe_load(A.values + idx);
        const Fr l0 = fe_load(A.l0 + idx), l_last = fe_load(A.l_last + idx), l_active = fe_load(A.l_active + idx);
        // l_0(X) * (1 - z_0(X))
        Fr z_first = fe_load(A.sets[0] + idx);
        value = fe_add(fe_mul(value, A.y), fe_mul(fe_sub(A.one, z_first), l0));
        // l_last(X) * (z_l(X)^2 - z_l(X))
        Fr z_l = fe_load(A.sets[A.n_sets - 1] + idx);
        value = fe_add(fe_mul(value, A.y), fe_mul(fe_sub(fe_sqr_w(z_l), z_l), l_last));
        // l_0(X) * (z_i(X) - z_{i-1}(omega^(last) X))
        for (uint32_t s = 1; s < A.n_sets; s++) {
            Fr t = fe_sub(fe_load(A.sets[s] + idx), fe_load(A.sets[s - 1] + r_last));
            value = fe_add(fe_mul(value, A.y), fe_mul(t, l0));
        }
        // (1 - (l_last + l_blind)) * (z_i(omega X) prod (p + beta s_j + gamma) - z_i(X) prod (p + delta^j beta X + gamma))
        Fr current_delta = fe_mul(A.delta_start, beta_term);
        for (uint32_t s = 0; s < A.n_sets; s++) {
            const uint32_t c0 = s * A.chunk_len;
            const uint32_t c1 = c0 + A.chunk_len < A.n_cols ? c0 + A.chunk_len : A.n_cols;
            Fr left = fe_load(A.sets[s] + r_next);
            for (uint32_t c = c0; c < c1; c++) {
                Fr v = fe_load(A.col_values[c] + idx);
                Fr t = fe_add(fe_add(v, fe_mul(A.beta, fe_load(A.col_perms[c] + idx))), A.gamma);
                left = fe_mul(left, t);
            }
            Fr right = fe_load(A.sets[s] + idx);
            for (uint32_t c = c0; c < c1; c++) {
                Fr v = fe_load(A.col_values[c] + idx);
                right = fe_mul(right, fe_add(fe_add(v, current_delta), A.gamma));
                current_delta = fe_mul(current_delta, A.delta);
            }
            value = fe_add(fe_mul(value, A.y), fe_mul(fe_sub(left, right), l_active));
        }
        fe_store(A.values + idx, value);
        beta_term = fe_mul(beta_term, stride);
    }
}

// evaluation.rs:517-565
__global__ void __launch_bounds__(128)
quotient_lookup_kernel(const LookupFoldArgs A) {
    const uint32_t nthreads = gridDim.x * blockDim.x;
    for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < A.size; idx += nthreads) {
        const uint32_t r_next = rotation_idx(idx, 1, A.rot_scale, A.size);
        const uint32_t r_prev = rotation_idx(idx, -1, A.rot_scale, A.size);
        Fr value = fe_load(A.values + idx);
        const Fr l0 = fe_load(A.l0 + idx), l_last = fe_load(A.l_last + idx), l_active = fe_load(A.l_active + idx);
        const Fr z = fe_load(A.product + idx), a_in = fe_load(A.permuted_input + idx), s_tab = fe_load(A.permuted_table + idx);
        const Fr a_minus_s = fe_sub(a_in, s_tab);
        value = fe_add(fe_mul(value, A.y), fe_mul(fe_sub(A.one, z), l0));
        value = fe_add(fe_mul(value, A.y), fe_mul(fe_sub(fe_sqr_w(z), z), l_last));
        Fr lhs = fe_mul(fe_mul(fe_load(A.product + r_next), fe_add(a_in, A.beta)), fe_add(s_tab, A.gamma));
        Fr rhs = fe_mul(z, fe_load(A.table_values + idx));
        value = fe_add(fe_mul(value, A.y), fe_mul(fe_sub(lhs, rhs), l_active));
        value = fe_add(fe_mul(value, A.y), fe_mul(a_minus_s, l0));
        Fr t = fe_mul(a_minus_s, fe_sub(a_in, fe_load(A.permuted_input + r_prev)));
        value = fe_add(fe_mul(value, A.y), fe_mul(t, l_active));
        fe_store(A.values + idx, value);
    }
}

// evaluation.rs:573-620
__global__ void __launch_bounds__(128)
quotient_shuffle_kernel(const ShuffleFoldArgs A) {
    const uint32_t nthreads = gridDim.x * blockDim.x;
    for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < A.size; idx += nthreads) {
        const uint32_t r_next = rotation_idx(idx, 1, A.rot_scale, A.size);
        Fr value = fe_load(A.values + idx);
        const Fr l0 = fe_load(A.l0 + idx), l_last = fe_load(A.l_last + idx), l_active = fe_load(A.l_active + idx);
        const Fr z = fe_load(A.product + idx);
        value = fe_add(fe_mul(value, A.y), fe_mul(fe_sub(A.one, z), l0));
        value = fe_add(fe_mul(value, A.y), fe_mul(fe_sub(fe_sqr_w(z), z), l_last));
        Fr t = fe_sub(fe_mul(fe_load(A.product + r_next), fe_load(A.shuffle_values + idx)), fe_mul(z, fe_load(A.input_values + idx)));
        value = fe_add(fe_mul(value, A.y), fe_mul(l_active, t));
        fe_store(A.values + idx, value);
    }
}
#endif  // __CUDACC__ && !H2B_NO_KERNELS

}  // namespace h2b
