// vr_session.h -- the one place where the C++ drop-in API touches the C-ABI (include/volrender.h).
//
// A Session owns one vr_context per CUDA device it drives.  Devices come from the environment
// variable VR_DEVICES ("0,1,2,3"; default "0").  Every integrator of the drop-in API flattens
// its Scene / Camera into the POD arrays of the C-ABI and calls through here; a non-zero status
// becomes a std::runtime_error carrying vr_last_error, which is the error behaviour of the
// reference's loaders and integrators (scene.h:47,79, gmm.h:637).  There is NO CPU fallback: with
// no usable device, Session::get() throws.
#pragma once
#include <cstdlib>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "volrender.h"

namespace vrhost {

class Session {
    std::vector<vr_context*> ctxs_;
    std::vector<int> devices_;
    bool comm_ready_ = false;

    Session() {
        const char* env = std::getenv("VR_DEVICES");
        std::string spec = env && *env ? env : "0";
        std::stringstream ss(spec);
        std::string tok;
        while (std::getline(ss, tok, ',')) if (!tok.empty()) devices_.push_back(std::atoi(tok.c_str()));
        if (devices_.empty()) devices_.push_back(0);
        for (int d : devices_) {
            vr_context* c = nullptr;
            const int rc = vr_create(&c, d);
            if (rc != VR_OK) {
                std::string why = c ? vr_last_error(c) : "no usable CUDA device";
                if (c) vr_destroy(c);
                for (vr_context* o : ctxs_) vr_destroy(o);
                ctxs_.clear();
                throw std::runtime_error("volrender_b200: cannot create a context on device " + std::to_string(d) + " (" + why +
                                         "); this renderer has no CPU fallback");
            }
            ctxs_.push_back(c);
        }
    }

public:
    Session(const Session&) = delete;
    Session& operator=(const Session&) = delete;
    ~Session() { for (vr_context* c : ctxs_) vr_destroy(c); }

    static Session& get() {
        static Session s;
        return s;
    }
    int num_devices() const { return int(ctxs_.size()); }
    int device_id(int i) const { return devices_[size_t(i)]; }
    vr_context* ctx(int i) const { return ctxs_[size_t(i)]; }

    // The inverse path's one collective (reference reduction inverse_integrator.h:157-198): the contexts of
    // this process form one NCCL communicator (ncclCommInitAll, made on first use) and every iteration sums the int64
    // partial F_diff of all devices in place, on the devices' own streams -- no copy to the host.
    void allreduce_sfd() {
        if (num_devices() < 2) return;
        if (!comm_ready_) {
            if (vr_comm_init_all(ctxs_.data(), num_devices()) != VR_OK) throw std::runtime_error(std::string("volrender_b200: ") + vr_last_error(ctxs_[0]));
            comm_ready_ = true;
        }
        if (vr_sfd_allreduce_all(ctxs_.data(), num_devices()) != VR_OK) throw std::runtime_error(std::string("volrender_b200: ") + vr_last_error(ctxs_[0]));
    }

    // status -> exception
    void check(int i, int rc) const {
        if (rc != VR_OK) throw std::runtime_error(std::string("volrender_b200: ") + vr_last_error(ctxs_[size_t(i)]));
    }
};

}  // namespace vrhost
