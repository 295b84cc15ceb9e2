/*
 * ros/ros.h — stub of the two ros::NodeHandle calls LeggedParam::load makes
 * (src/legged_ctrl/src/LeggedState.cpp:20-209): getParam / param, backed by a flat name -> number
 * table that ros::shim::load_yaml fills from one of the reference's config/*.yaml files (the
 * launch files load them into the global namespace: launch/gazebo_a1_convex.launch:7).
 * TEST INFRASTRUCTURE ONLY; see mini_eigen.h.
 */
#ifndef LMPC_SHIM_ROS_H_
#define LMPC_SHIM_ROS_H_
#include <cstdlib>
#include <fstream>
#include <map>
#include <string>

namespace ros {
namespace shim {
inline std::map<std::string, double>& table() { static std::map<std::string, double> t; return t; }
inline std::string key(const std::string& name) { return (!name.empty() && name[0] == '/') ? name.substr(1) : name; }
/* flat `name: number  # comment` lines; anything else is ignored */
inline int load_yaml(const char* path) {
  std::ifstream f(path);
  if (!f) return -1;
  table().clear();
  std::string line; int n = 0;
  while (std::getline(f, line)) {
    size_t h = line.find('#'); if (h != std::string::npos) line.erase(h);
    size_t c = line.find(':'); if (c == std::string::npos) continue;
    std::string k = line.substr(0, c), v = line.substr(c + 1);
    size_t a = k.find_first_not_of(" \t"), b = k.find_last_not_of(" \t");
    if (a == std::string::npos) continue;
    k = k.substr(a, b - a + 1);
    a = v.find_first_not_of(" \t"); if (a == std::string::npos) continue;
    v = v.substr(a);
    double x;
    if (v.compare(0, 4, "true") == 0) x = 1; else if (v.compare(0, 5, "false") == 0) x = 0;
    else { char* end = nullptr; x = std::strtod(v.c_str(), &end); if (end == v.c_str()) continue; }
    table()[k] = x; ++n;
  }
  return n;
}
}  // namespace shim

class NodeHandle {
 public:
  NodeHandle() {}
  template <class T>
  bool getParam(const std::string& name, T& val) const {
    std::map<std::string, double>::const_iterator it = shim::table().find(shim::key(name));
    if (it == shim::table().end()) return false;
    val = (T)it->second; return true;
  }
  template <class T>
  bool param(const std::string& name, T& val, const T& def) const {
    if (getParam(name, val)) return true;
    val = def; return false;
  }
};
inline bool ok() { return true; }
}  // namespace ros
#endif
