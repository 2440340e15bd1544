// xml_frontend.cpp -- TEST PROGRAM for iemic_b200/host/ParameterListXML.H (host only, no GPU): reads an ocean parameter file of
// the reference and prints the configuration, the starting parameters and the mask the C++ host mirror would hand to
// thcm_b200_create; tests/test_params.py compares the lines with iemic_b200/params.py and iemic_b200/checkpoint.py.
//   xml_frontend <ocean_params.xml> [<mask dir> <mask out.bin>]      exit code 3 + message when the file is rejected
//   xml_frontend --list <any.xml>                                    prints every parameter of the file, nested
#include <cstdio>
#include <string>
#include "ParameterListXML.H"

using namespace iemic_b200;

static void show(const ParameterList& p, const std::string& prefix) {
  for (auto& kv : p.values()) {
    std::printf("%s%s = ", prefix.c_str(), kv.first.c_str());
    const ParameterValue& v = kv.second;
    if (v.is<bool>()) std::printf("bool %d\n", (int)v.as<bool>());
    else if (v.is<int>()) std::printf("int %d\n", v.as<int>());
    else if (v.is<double>()) std::printf("double %.17g\n", v.as<double>());
    else if (v.is<char>()) std::printf("char %c\n", v.as<char>());
    else std::printf("string %s\n", v.as<std::string>().c_str());
  }
  for (auto& kv : p.sublists()) show(kv.second, prefix + kv.first + "/");
}

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  try {
    if (std::string(argv[1]) == "--list") { show(readParameterListXML(argv[2]), ""); return 0; }
    const ThcmSetup s = thcmConfig(readParameterListXML(argv[1]));
    const thcm_b200_config& c = s.cfg;
    std::printf("cfg n %d\ncfg m %d\ncfg l %d\ncfg xmin %.17g\ncfg xmax %.17g\ncfg ymin %.17g\ncfg ymax %.17g\n", c.n, c.m, c.l, c.xmin, c.xmax, c.ymin, c.ymax);
    std::printf("cfg periodic %d\ncfg hdim %.17g\ncfg qz %.17g\ncfg alphaT %.17g\ncfg alphaS %.17g\ncfg ih %d\ncfg vmix %d\ncfg tap %d\n", c.periodic, c.hdim, c.qz, c.alphaT, c.alphaS, c.ih, c.vmix, c.tap);
    std::printf("cfg rho_mixing %d\ncfg coriolis_on %d\ncfg forcing_type %d\ncfg tres %d\ncfg sres %d\ncfg iza %d\ncfg ite %d\ncfg its %d\n", c.rho_mixing, c.coriolis_on, c.forcing_type, c.tres, c.sres, c.iza, c.ite, c.its);
    std::printf("cfg coupled_T %d\ncfg coupled_S %d\ncfg int_sign %d\ncfg intcond_i %d\ncfg intcond_j %d\ncfg fix_pressure_points %d\n", c.coupled_T, c.coupled_S, c.int_sign, c.intcond_i, c.intcond_j, c.fix_pressure_points);
    for (auto& kv : s.startingParameters) std::printf("start %s = %.17g\n", kv.first.c_str(), kv.second);
    std::printf("mask %s\n", s.maskFile.c_str());
    if (argc > 3 && !s.maskFile.empty()) {
      const std::vector<int> mask = readMaskFile(std::string(argv[2]) + "/" + s.maskFile, c.n, c.m, c.l);
      FILE* f = std::fopen(argv[3], "wb");
      if (!f) return 2;
      std::fwrite(mask.data(), sizeof(int), mask.size(), f);
      std::fclose(f);
    }
  } catch (const std::exception& e) {
    std::printf("rejected: %s\n", e.what());
    return 3;
  }
  return 0;
}
