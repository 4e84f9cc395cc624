// optimizer.h -- AdamOptimizer of the drop-in API (reference optimizer.h:13-57).
// Same constructor, step() contract (false on a size mismatch, parameters updated in place) and
// reset_state()/dim(); the update itself is the device kernel behind vr_adam_step, which rounds
// exactly like the reference's float expression (tests/test_gpu_parity.py::test_adam_step_bit_exact).
#pragma once
#include <algorithm>
#include <vector>

#include "vr_session.h"

class AdamOptimizer {
    std::vector<float> m, v;
    float lr, beta1, beta2, eps;
    int t;

public:
    AdamOptimizer(size_t ndim, float lr_in = 1e-3f, float beta1_in = 0.9f, float beta2_in = 0.999f, float eps_in = 1e-8f)
        : m(ndim, 0.0f), v(ndim, 0.0f), lr(lr_in), beta1(beta1_in), beta2(beta2_in), eps(eps_in), t(0) {}

    bool step(std::vector<float>& params, const std::vector<float>& grads) {
        if (params.size() != grads.size() || params.size() != m.size()) return false;
        if (params.empty()) { ++t; return true; }
        vrhost::Session& S = vrhost::Session::get();
        S.check(0, vr_adam_step(S.ctx(0), int(params.size()), params.data(), grads.data(), m.data(), v.data(), &t, lr, beta1, beta2, eps));
        return true;
    }
    void reset_state() {
        std::fill(m.begin(), m.end(), 0.0f);
        std::fill(v.begin(), v.end(), 0.0f);
        t = 0;
    }
    size_t dim() const { return m.size(); }
    int steps_taken() const { return t; }
};
