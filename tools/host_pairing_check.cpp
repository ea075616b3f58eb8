// Host-side check and timing of the engine's pairing module (kateth_b200/csrc/host_bls.cpp), no GPU needed:
//   g++ -O3 -march=x86-64-v3 -std=c++17 -DKB200_HOST_SELFTEST -I kateth_b200/csrc tools/host_pairing_check.cpp kateth_b200/csrc/host_bls.cpp -o /tmp/hpc && /tmp/hpc
// prints whether the Granger-Scott cyclotomic squaring equals the generic Fp12 squaring on an element of the cyclotomic subgroup,
// whether e(-G1, G2) e(G1, G2) == 1 and e(-G1, G2) e(2 G1, G2) != 1, and the time of one two-pairing check.  tests/test_host_bls.py runs it.
#include "host_bls.hpp"
#include <chrono>
#include <cstdio>
using namespace kb200::host;
namespace kb200 { namespace host { bool selftest_cyclotomic(const G1A&, const G2Prepared&); double selftest_fmul_ns(int); } }
static void parse(const char* h, uint32_t* l) { for (int i = 0; i < 12; i++) { unsigned v; sscanf(h + 8 * (11 - i), "%8x", &v); l[i] = v; } }
static G1A point(const char* x, const char* y) {
    uint32_t lx[12], ly[12]; parse(x, lx); parse(y, ly);
    G1A a; a.inf = false; fp_from_plain_u32(a.x, lx); fp_from_plain_u32(a.y, ly); return a;
}
int main() {
    G2A g; g2_generator(g); G2Prepared p; g2_prepare(p, g);
    const G1A g1 = point("17f1d3a73197d7942695638c4fa9ac0fc3688c4f9774b905a14e3a3f171bac586c55e83ff97a1aeffb3af00adb22c6bb",
                         "08b3f481e3aaa0f1a09e30ed741d8ae4fcf5e095d5d00af600db18cb2c04b3edd03cc744a2888ae40caa232946c5e7e1");
    const G1A g1x2 = point("0572cbea904d67468808c8eb50a9450c9721db309128012543902d0ac358a62ae28f75bb8f1c7c42c39a8c5529bf0f4e",
                           "166a9d8cabc673a322fda673779d8e3822ba3ecb8670e461f73bb9021d5fd76a4c56d9d4cd16bd1bba86881979749d28");
    const bool cyc = selftest_cyclotomic(g1, p), same = verify_pairings(g1, p, g1, p), diff = verify_pairings(g1, p, g1x2, p);
    printf("cyclotomic_equals_generic %d\npairing_same %d\npairing_different %d\n", (int)cyc, (int)same, (int)diff);
    const int N = 50;
    auto t0 = std::chrono::steady_clock::now();
    bool ok = true;
    for (int i = 0; i < N; i++) ok = verify_pairings(g1, p, g1, p) && ok;
    printf("fmul_ns %.1f\n", selftest_fmul_ns(2000000));
    printf("ms_per_check %.3f\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() / N);
    return (cyc && same && !diff && ok) ? 0 : 1;
}
