// oracle/shim: TEST INFRASTRUCTURE ONLY. badger_file_lib (not in the checkout): a plain ofstream.
#pragma once
#include <fstream>
#include <string>
namespace badger_file_lib
{
class atomic_ofstream : public std::ofstream
{
public:
  explicit atomic_ofstream(const std::string& path) : std::ofstream(path) {}
};
}  // namespace badger_file_lib
