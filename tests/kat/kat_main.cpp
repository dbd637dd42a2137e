// kat_main.cpp — runner: kat_runner --lib <path> --prefix <oe_|oeo_> [--list] [--test NAME]...
#include <cstring>

#include "kat.hpp"

int main(int argc, char** argv) {
  std::string lib, prefix = "oe_";
  std::vector<std::string> only;
  bool list = false;
  for (int k = 1; k < argc; ++k) {
    if (!std::strcmp(argv[k], "--lib") && k + 1 < argc) lib = argv[++k];
    else if (!std::strcmp(argv[k], "--prefix") && k + 1 < argc) prefix = argv[++k];
    else if (!std::strcmp(argv[k], "--test") && k + 1 < argc) only.push_back(argv[++k]);
    else if (!std::strcmp(argv[k], "--list")) list = true;
  }
  if (list) {
    for (auto& [name, fn] : kat::registry()) std::cout << name << "\n";
    return 0;
  }
  if (lib.empty()) { std::cerr << "usage: kat_runner --lib <path> --prefix <oe_|oeo_> [--list] [--test NAME]\n"; return 2; }
  std::unique_ptr<sdk::EngineLibrary> library;
  if (prefix == "oe_") library.reset(new sdk::EngineLibrary(sdk::EngineLibrary::open(lib)));
  else library.reset(new kat::OracleLibrary(lib, prefix));
  kat::library() = library.get();
  int failed = 0, ran = 0;
  for (auto& [name, fn] : kat::registry()) {
    if (!only.empty() && std::find(only.begin(), only.end(), name) == only.end()) continue;
    ++ran;
    try {
      fn();
      std::cout << "PASS " << name << std::endl;
    } catch (const std::exception& e) {
      ++failed;
      std::cout << "FAIL " << name << ": " << e.what() << std::endl;
    }
  }
  std::cout << ran - failed << " passed, " << failed << " failed" << std::endl;
  return failed ? 1 : (ran ? 0 : 3);
}
