// sirrt_validators -- drives MDP::StateValidityCheckerFunctor / MDP::SpaceTimeMotionValidator (OMPLValidators.hpp)
// over motions read from a text file; tests/test_gpu_parity.py compares its output with the oracle's
// restatement of SpaceTimeMotionValidator.cpp:22-127.
// usage: sirrt_validators <scene.pack> <motions.txt> [device]
//   motions.txt: one motion per line: dof values of s1, t1 [s], dof values of s2, t2 [s]
// output per motion: "<valid 2-arg form> <valid lastValid form> <lastValid fraction> <isValid(s2)> <clearance(s1)>"
// followed by one line "batch <code> <code> ..." from checkMotions (variant 0) and the validation counters.
#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>

#include "OMPLValidators.hpp"

int main(int argc, char** argv) {
  if (argc < 3) { std::fprintf(stderr, "usage: %s scene.pack motions.txt [device]\n", argv[0]); return 2; }
  try {
    std::ifstream pf(argv[1], std::ios::binary);
    if (!pf) throw std::runtime_error("cannot open scene pack");
    std::vector<unsigned char> pack((std::istreambuf_iterator<char>(pf)), std::istreambuf_iterator<char>());
    MDP::StateValidityCheckerFunctor svc(pack, argc > 3 ? std::atoi(argv[3]) : 0);
    MDP::SpaceTimeMotionValidator mv(svc);
    const int dof = svc.manager().dof();
    std::ifstream mf(argv[2]);
    std::vector<MDP::SpaceTimeState> from, to;
    std::string line;
    while (std::getline(mf, line)) {
      std::istringstream is(line);
      MDP::SpaceTimeState a, b;
      a.values.resize(dof); b.values.resize(dof);
      bool ok = true;
      for (int i = 0; i < dof; i++) ok = ok && (is >> a.values[i]);
      ok = ok && (is >> a.position);
      for (int i = 0; i < dof; i++) ok = ok && (is >> b.values[i]);
      ok = ok && (is >> b.position);
      if (!ok) continue;
      from.push_back(a); to.push_back(b);
    }
    for (size_t e = 0; e < from.size(); e++) {
      MDP::SpaceTimeState last;
      std::pair<MDP::SpaceTimeState*, double> lv{&last, -1.0};
      const bool v0 = mv.checkMotion(&from[e], &to[e]);
      const bool v1 = mv.checkMotion(&from[e], &to[e], lv);
      std::printf("%d %d %.17g %d %.9g\n", (int)v0, (int)v1, lv.second, (int)svc.isValid(&to[e]), svc.clearance(&from[e]));
    }
    std::vector<int32_t> codes;
    mv.checkMotions(from, to, 0, codes);
    std::printf("batch");
    for (int32_t c : codes) std::printf(" %d", c);
    std::printf("\ncounters %d %d\n", svc.get_count_of_validation(), mv.get_count_of_validation());
  } catch (const std::exception& e) {
    std::fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
  return 0;
}
