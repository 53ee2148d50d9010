// ModelArray / ModelState / ModelComponent facade test.
//   test_modelarray --cpu   layout, transposition and registry logic only (no GPU needed)
//   test_modelarray         additionally drives GpuDynamics / GpuColumnPhysics on cuda:0 and
//                           checks them bit-for-bit against the same run through the bare C ABI
// Needs no reference headers, so it builds everywhere (host/Makefile target test_modelarray).
#include "GpuComponents.hpp"

#include <cmath>
#include <cstdio>
#include <cstring>

using namespace Nextsim;

static int failures = 0;
#define CHECK(cond)                                                                               \
    do {                                                                                          \
        if (!(cond)) {                                                                            \
            printf("CHECK failed at line %d: %s\n", __LINE__, #cond);                             \
            ++failures;                                                                           \
        }                                                                                         \
    } while (0)

class Dummy : public ModelComponent {
public:
    Dummy() { registerModule(); }
    std::string getName() const override { return "Dummy"; }
    void setData(const ModelState& s) override
    {
        auto it = s.find("x");
        if (it != s.end())
            x = it->second;
    }
    ModelState getState() const override { return { { "x", x } }; }
    ModelArray x = ModelArray::HField();
};

static void cpu_tests()
{
    ModelArray::setDimensions(5, 7);
    ModelArray::setNComponents(ModelArray::Type::DG, 6);
    ModelArray h = ModelArray::HField(), d = ModelArray::DGField(), s = ModelArray::DGStressField();
    CHECK(h.size() == 35 && d.size() == 35 * 6 && s.size() == 35 * 3 && d.trueSize() == 35);
    for (size_t j = 0; j < 5; ++j)
        for (size_t i = 0; i < 7; ++i)
            for (size_t c = 0; c < 6; ++c)
                d(j, i, c) = 100. * j + 10. * i + c;
    CHECK(d[(2 * 7 + 3) * 6 + 4] == 234.); // component fastest, as upstream
    std::vector<double> planes = d.toComponentMajor();
    CHECK(planes[4 * 35 + 2 * 7 + 3] == 234.); // [component][row][column], as the device
    ModelArray e = ModelArray::DGField();
    e.fromComponentMajor(planes);
    CHECK(std::memcmp(e.data(), d.data(), d.size() * sizeof(double)) == 0);
    bool threw = false;
    try {
        e.fromComponentMajor(std::vector<double>(3));
    } catch (const std::length_error&) {
        threw = true;
    }
    CHECK(threw);
    {
        Dummy a;
        CHECK(ModelComponent::nRegistered() == 1);
        ModelArray x = ModelArray::HField();
        x = 3.5;
        ModelComponent::setAllModuleData({ { "x", x }, { "unrelated", h } });
        ModelState all = ModelComponent::getAllModuleStates();
        CHECK(all.count("x") == 1 && all.at("x")[34] == 3.5);
    }
    CHECK(ModelComponent::nRegistered() == 0);
}

static double rnd(unsigned long long& s)
{
    s += 0x9E3779B97F4A7C15ull;
    unsigned long long z = s;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (z >> 11) * (1.0 / 9007199254740992.0);
}

static void gpu_tests()
{
    const size_t ny = 40, nx = 96;
    const double hx = 2000., hy = 2000.;
    ModelArray::setDimensions(ny, nx);
    ModelArray::setNComponents(ModelArray::Type::DG, 6);
    unsigned long long seed = 7;
    ModelState st;
    for (const char* nm : { "hice", "cice" }) {
        ModelArray a = ModelArray::DGField();
        for (size_t p = 0; p < ny * nx; ++p) {
            a[p * 6] = nm[0] == 'h' ? 0.3 + 0.1 * rnd(seed) : 0.8 + 0.2 * rnd(seed);
            for (int c = 1; c < 6; ++c)
                a[p * 6 + c] = 0.01 * (rnd(seed) - 0.5);
        }
        st.emplace(nm, a);
    }
    for (const char* nm : { "u", "v", "uwind", "vwind", "uocean", "vocean" }) {
        ModelArray a = ModelArray::VertexField();
        const double sc = nm[1] == 'w' ? 10. : 0.05;
        for (size_t p = 0; p < a.size(); ++p)
            a[p] = sc * (rnd(seed) - 0.5);
        st.emplace(nm, a);
    }
    for (const char* nm : { "s11", "s12", "s22" }) {
        ModelArray a = ModelArray::DGStressField();
        for (size_t p = 0; p < a.size(); ++p)
            a[p] = 50. * (rnd(seed) - 0.5);
        st.emplace(nm, a);
    }
    GpuContext gpu(0);
    {
        GpuDynamics dyn(gpu, hx, hy);
        ModelComponent::setAllModuleData(st);
        dyn.update(120., 7);
        dyn.update(120., 7);
        ModelState got = ModelComponent::getAllModuleStates();
        CHECK(got.size() == 7);
        // the same through the bare C ABI on a second context
        GpuContext ref(0);
        nsdg_dyn_config cfg = { (int64_t)nx, (int64_t)ny, (int64_t)ny, 0, hx, hy, 6, 0, 0, 0 };
        ref.check(nsdg_dyn_init(ref.ctx, &cfg), "init");
        for (const auto& f : GpuDynamics::fields()) {
            std::vector<double> planes = st.at(f.first).toComponentMajor();
            ref.check(nsdg_dyn_upload(ref.ctx, f.second, planes.data()), "upload");
        }
        ref.check(nsdg_dyn_step(ref.ctx, 120., 7), "step");
        ref.check(nsdg_dyn_step(ref.ctx, 120., 7), "step");
        for (const auto& f : GpuDynamics::fields()) {
            if (f.second >= NSDG_DF_UATM)
                continue;
            const ModelArray& a = got.at(f.first);
            std::vector<double> planes(a.size());
            ref.check(nsdg_dyn_download(ref.ctx, f.second, planes.data()), "download");
            ModelArray b(a.getType());
            b.fromComponentMajor(planes);
            CHECK(std::memcmp(a.data(), b.data(), a.size() * sizeof(double)) == 0);
            double mx = 0;
            for (size_t p = 0; p < a.size(); ++p)
                mx = std::fmax(mx, std::fabs(a[p]));
            CHECK(std::isfinite(mx) && mx > 0);
        }
        printf("GpuDynamics component == C ABI (7 fields, bitwise)\n");
    }
    {
        GpuColumnPhysics phys(gpu);
        ModelState s;
        auto H = [&](const char* nm, double lo, double hi) {
            ModelArray a = ModelArray::HField();
            for (size_t p = 0; p < a.size(); ++p)
                a[p] = lo + (hi - lo) * rnd(seed);
            s.emplace(nm, a);
        };
        H("cice_mean", 0.3, 1.);
        H("hice_mean", 0.1, 2.);
        H("hsnow", 0., 0.2);
        H("sst", -1.8, 1.);
        H("sss", 30., 35.);
        H("tice", -15., -1.);
        H("tair", -20., 2.);
        H("tdew", -25., -3.);
        H("slp", 99000., 102000.);
        H("qsw", 0., 200.);
        H("qlw", 200., 320.);
        H("mld", 10., 40.);
        H("snowfall", 0., 1e-4);
        H("wind", 0., 15.);
        phys.setData(s);
        phys.update(600.);
        ModelState got = phys.getState();
        const ModelArray& c = got.at("cice_mean");
        size_t changed = 0;
        for (size_t p = 0; p < c.size(); ++p) {
            CHECK(std::isfinite(c[p]) && c[p] >= 0. && c[p] <= 1.5);
            changed += c[p] != s.at("cice_mean")[p];
        }
        CHECK(changed > c.size() / 2);
        CHECK(std::memcmp(got.at("sst").data(), s.at("sst").data(), c.size() * sizeof(double)) == 0); // Q5
        printf("GpuColumnPhysics component stepped %zu cells\n", c.size());
    }
}

int main(int argc, char** argv)
{
    cpu_tests();
    if (!(argc > 1 && std::strcmp(argv[1], "--cpu") == 0)) {
        try {
            gpu_tests();
        } catch (const std::exception& e) {
            printf("exception: %s\n", e.what());
            ++failures;
        }
    }
    printf(failures ? "FAIL\n" : "PASS\n");
    return failures ? 1 : 0;
}
