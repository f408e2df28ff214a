// Host-only check of turtle::sc::json (the serializer's document type): prints dump(4) and dump().
#include <iomanip>
#include <iostream>

#include "sea_current_b200.hpp"

int main() {
  turtle::sc::json j;
  j.samples.push_back({0.0, 0.0, 0.5, 0.0, 0.0, 0.0});
  j.samples.push_back({(double)0.02f, (double)0.1f, -0.25, 1e-7, 10.5, -3.0});
  j.samples.push_back({86.5553, 0.2531, 1.0 / 3.0, -2.5e10, 1e-5, 123456789.0});
  std::cout << std::setw(4) << j << std::endl;
  std::cout << "====" << std::endl;
  std::cout << j << std::endl;
  std::cout << "====" << std::endl;
  std::cout << turtle::sc::json().dump(4) << std::endl;
  return 0;
}
