// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or shipped with the product path.
//
// pybind11 stand-in for the reference's Boost.Python module `pyevstation`
// (reference: evcssp_env_cpp/envs/lion_cpp20/main.cpp:19-290).  Boost.Python cannot be built in
// this image (Boost core headers absent, no network), so this file binds the UNMODIFIED reference
// header SCP_Base/CHS.hpp -- included from where it lies under /root/reference, never copied --
// and exposes exactly the names the reference's Python layer imports:
//   Aggregator_Simple.py:7  SlowChargeStation, FastChargeStation, Vector_float
//   hydro_sys.py:6          PoissonNumber, CarArriveRandom, RandomUtil
//   evcssp_manager.py:16    Change_Use_Seed
// plus test hooks that make the two process-global random streams reproducible per hub:
//   * the U stream: CHS.hpp:41 calls libc rand(); this object defines its own rand() (bound with
//     -Wl,-Bsymbolic so CHS.hpp's calls resolve here) backed by the counter hash that the CUDA
//     engine and oracle/hub_oracle.c also use (see include/chub_rng.h);
//   * the E stream: `std::default_random_engine e` is a non-static global (CHS.hpp:25) so the
//     hook simply calls e.seed().
// Build recipe: oracle/Makefile (outputs only into oracle/_ref/).
#include <fstream>
#include <string>
#include <vector>
#include <algorithm>
#include <random>
#include <cstdlib>
#include <ctime>
#include <map>
#include <cstdint>
#include <sstream>

#include "../include/chub_rng.h"

static uint32_t g_u_key0 = 0, g_u_key1 = 0, g_u_ctr = 0;
static bool g_u_hooked = false;
static unsigned long long g_u_calls = 0;

// When not hooked fall back to a private glibc-TYPE_3-free LCG so behaviour stays deterministic.
extern "C" int rand(void) noexcept {
    ++g_u_calls;
    if (g_u_hooked) {
        return (int) chub_urand(g_u_key0, g_u_key1, g_u_ctr++);
    }
    static uint32_t s = 12345u;
    s = s * 1103515245u + 12345u;
    return (int) ((s >> 1) & 0x7fffffffu);
}

#include "SCP_Base/CHS.hpp"

#include <pybind11/pybind11.h>
#include <pybind11/stl.h>
#include <pybind11/stl_bind.h>
#include <pybind11/numpy.h>

namespace py = pybind11;

PYBIND11_MAKE_OPAQUE(std::vector<float>);

template <class StationT>
static py::dict dump_station(StationT &st, const char *prefix) {
    const int n = st.charge_number;
    py::array_t<float> car(n), charge(n), emergency(n), power(n), soc(n), init_soc(n), target(n);
    py::array_t<float> p_arrive(n), p_target(n), p_soc(n);
    py::array_t<int> stay(n), already(n), must(n);
    for (int i = 0; i < n; ++i) {
        car.mutable_at(i) = st.situation["car"][i];
        charge.mutable_at(i) = st.situation["charge"][i];
        emergency.mutable_at(i) = st.situation["emergency"][i];
        power.mutable_at(i) = st.situation["power"][i];
        soc.mutable_at(i) = st.situation["soc"][i];
        init_soc.mutable_at(i) = st.situation["init_soc"][i];
        target.mutable_at(i) = st.situation["target_soc"][i];
        auto &p = st.positions.at(std::string(prefix) + std::to_string(i));
        p_arrive.mutable_at(i) = p.arrive_soc;
        p_target.mutable_at(i) = p.target_soc;
        p_soc.mutable_at(i) = p.current_soc;
        stay.mutable_at(i) = p.stay_time;
        already.mutable_at(i) = p.already_stay_time;
        must.mutable_at(i) = p.must_charge ? 1 : 0;
    }
    py::dict d;
    d["car"] = car; d["charge"] = charge; d["emergency"] = emergency; d["power"] = power;
    d["soc"] = soc; d["init_soc"] = init_soc; d["target_soc"] = target;
    d["pile_arrive_soc"] = p_arrive; d["pile_target_soc"] = p_target; d["pile_current_soc"] = p_soc;
    d["stay_time"] = stay; d["already_stay_time"] = already; d["must_charge"] = must;
    d["line"] = st.line; d["min_power"] = st.min_power; d["max_power"] = st.max_power;
    d["charge_power"] = st.charge_power; d["car_number"] = st.car_number;
    d["flow_in_last"] = st.flow_in_number.empty() ? -1 : st.flow_in_number.back();
    d["station_time_hole"] = st.station_time_hole;
    return d;
}

template <class StationT>
static void bind_station(py::module_ &m, const char *name, const char *prefix) {
    py::class_<StationT>(m, name)
        .def(py::init<int>())
        .def(py::init<int, bool>())
        .def(py::init<int, bool, bool>())
        .def("evs_step", [](StationT &s, std::vector<float> a) { s.evs_step(std::move(a), false); })
        .def("evs_step", [](StationT &s, float a) { s.evs_step(a); })
        .def("evs_reset", &StationT::evs_reset)
        .def("calculate_output", &StationT::calculate_output)
        .def("print_situation", &StationT::print_situation)
        .def("dump", [prefix](StationT &s) { return dump_station(s, prefix); })
        .def_readwrite("transformer_limit", &StationT::transformer_limit)
        .def_readwrite("constant_power", &StationT::constant_power)
        .def_readwrite("charge_number", &StationT::charge_number)
        .def_readwrite("min_power", &StationT::min_power)
        .def_readwrite("max_power", &StationT::max_power)
        .def_readwrite("charge_power", &StationT::charge_power)
        .def_readwrite("car_number", &StationT::car_number)
        .def_readwrite("line", &StationT::line)
        .def_readwrite("station_time_hole", &StationT::station_time_hole)
        .def_readwrite("constant_charging", &StationT::constant_charging)
        .def_readwrite("wait", &StationT::wait)
        .def_readwrite("max_line", &StationT::max_line)
        .def_property_readonly("flow_in_number", [](StationT &s) { return s.flow_in_number; });
}

PYBIND11_MODULE(pyevstation, m) {
    m.doc() = "test-only pybind11 binding of the unmodified reference CHS.hpp";
    py::bind_vector<std::vector<float>>(m, "Vector_float");
    m.def("Change_Use_Seed", &Change_Use_Seed);

    py::class_<RandomUtil>(m, "RandomUtil")
        .def(py::init<>())
        .def_static("uniform_rand", [](float a, float b) { return RandomUtil::uniform_rand(a, b); })
        .def_static("uniform_rand", [](float a, float b, bool s) { return RandomUtil::uniform_rand(a, b, s); });
    py::class_<PoissonNumber>(m, "PoissonNumber")
        .def(py::init<>())
        .def_static("give_car_number_wrt_poisson", &PoissonNumber::give_car_number_wrt_poisson)
        .def_static("ev_car_number_wrt_poisson_fast", &PoissonNumber::ev_car_number_wrt_poisson_fast)
        .def_static("ev_car_number_wrt_poisson_slow", &PoissonNumber::ev_car_number_wrt_poisson_slow)
        .def_static("hv_car_number_wrt_poisson", &PoissonNumber::hv_car_number_wrt_poisson);
    py::class_<CarArriveRandom>(m, "CarArriveRandom")
        .def(py::init<>())
        .def_static("mk_soc", &CarArriveRandom::mk_soc)
        .def_static("mk_late_time", &CarArriveRandom::mk_late_time)
        .def_static("init_station_car_number", [](int mu) { return CarArriveRandom::init_station_car_number(mu, 3); });

    // curve functions, for the exhaustive-ish curve parity tests of the restatement
    m.def("slow_t2p", [](float t, bool c) { return UtilSlow::slow_time_to_power(t, c); });
    m.def("slow_t2s", [](float t, bool c) { return UtilSlow::slow_time_to_soc(t, c); });
    m.def("slow_s2t", [](float s, bool c) { return UtilSlow::slow_soc_to_time(s, c); });
    m.def("fast_t2p", [](float t, bool c) { return UtilFast::fast_time_to_power(t, c); });
    m.def("fast_t2s", [](float t, bool c) { return UtilFast::fast_time_to_soc(t, c); });
    m.def("fast_s2t", [](float s, bool c) { return UtilFast::fast_soc_to_time(s, c); });
    m.def("curve_batch", [](int which, py::array_t<float> x, bool c) {
        auto in = x.unchecked<1>();
        py::array_t<float> out(in.shape(0));
        auto o = out.mutable_unchecked<1>();
        for (py::ssize_t i = 0; i < in.shape(0); ++i) {
            float v = in(i), r = 0;
            switch (which) {
                case 0: r = UtilSlow::slow_time_to_power(v, c); break;
                case 1: r = UtilSlow::slow_time_to_soc(v, c); break;
                case 2: r = UtilSlow::slow_soc_to_time(v, c); break;
                case 3: r = UtilFast::fast_time_to_power(v, c); break;
                case 4: r = UtilFast::fast_time_to_soc(v, c); break;
                case 5: r = UtilFast::fast_soc_to_time(v, c); break;
            }
            o(i) = r;
        }
        return out;
    });

    bind_station<SlowChargeStation>(m, "SlowChargeStation", "SP");
    bind_station<FastChargeStation>(m, "FastChargeStation", "FP");

    // ---- test hooks (not part of the reference surface) ----
    m.def("hook_seed", [](uint32_t k0, uint32_t k1, uint32_t ctr, uint32_t e_state) {
        g_u_key0 = k0; g_u_key1 = k1; g_u_ctr = ctr; g_u_hooked = true;
        e.seed(e_state);
    }, "seed both per-hub streams: U counter hash keys/counter and minstd_rand0 state");
    m.def("hook_off", []() { g_u_hooked = false; });
    m.def("hook_u_counter", []() { return g_u_ctr; });
    m.def("hook_u_calls", []() { return g_u_calls; });
    m.def("hook_e_state", []() {  // minstd_rand0 prints its state as a decimal integer
        std::ostringstream os; os << e; return (uint32_t) std::stoul(os.str());
    });
    m.def("flow_table", []() {
        const size_t r = car_flow_possibility_list.size();
        const size_t c = r ? car_flow_possibility_list[0].size() : 0;
        py::array_t<double> out({r, c});
        auto o = out.mutable_unchecked<2>();
        for (size_t i = 0; i < r; ++i)
            for (size_t j = 0; j < c; ++j) o(i, j) = car_flow_possibility_list[i][j];
        return out;
    });
    m.def("set_flow_table", [](py::array_t<double> t) {
        auto a = t.unchecked<2>();
        car_flow_possibility_list.assign(a.shape(0), std::vector<double>(a.shape(1)));
        for (py::ssize_t i = 0; i < a.shape(0); ++i)
            for (py::ssize_t j = 0; j < a.shape(1); ++j) car_flow_possibility_list[i][j] = a(i, j);
    }, "replace the arrival CDF table (the GPU box has no CSV in the CWD)");
}
