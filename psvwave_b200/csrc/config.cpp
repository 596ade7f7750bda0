#include "config.hpp"

#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <iostream>

namespace psv {

namespace {

std::string rstrip(std::string s) {
  while (!s.empty() && std::isspace(static_cast<unsigned char>(s.back()))) s.pop_back();
  return s;
}

std::string lstrip(const std::string &s) {
  size_t i = 0;
  while (i < s.size() && std::isspace(static_cast<unsigned char>(s[i]))) ++i;
  return s.substr(i);
}

// Cut an inline comment: a ';' counts only when the previous character is whitespace
// (ext/inih/INIReader.h:140-154).
std::string cut_inline_comment(const std::string &s) {
  bool prev_space = false;
  for (size_t i = 0; i < s.size(); ++i) {
    if (prev_space && s[i] == ';') return s.substr(0, i);
    prev_space = std::isspace(static_cast<unsigned char>(s[i])) != 0;
  }
  return s;
}

std::string lower(std::string s) {
  std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return std::tolower(c); });
  return s;
}

void strip_all_whitespace(std::string &s) {
  s.erase(std::remove_if(s.begin(), s.end(), [](unsigned char c) { return std::isspace(c); }),
          s.end());
}

std::vector<std::string> split(const std::string &s, const std::string &delim) {
  std::vector<std::string> out;
  size_t from = 0, pos;
  while ((pos = s.find(delim, from)) != std::string::npos) {
    out.push_back(s.substr(from, pos - from));
    from = pos + delim.size();
  }
  out.push_back(s.substr(from));
  return out;
}

struct IniValues {
  std::map<std::string, std::string> values;
  std::string path;

  const std::string &get(const std::string &section, const std::string &name) const {
    const std::string key = lower(section + "=" + name);
    auto it = values.find(key);
    if (it == values.end())
      throw invalid_argument("The requested key (" + key + ") does not exist in the .ini file (" +
                             path + ").");
    return it->second;
  }
  long get_integer(const std::string &section, const std::string &name) const {
    const std::string &v = get(section, name);
    char *end = nullptr;
    const long n = std::strtol(v.c_str(), &end, 0);  // decimal, 0x hex (INIReader.h:417-429)
    if (end == v.c_str())
      throw invalid_argument("The requested key (" + lower(section + "=" + name) +
                             ") does not seem to be an integer in the .ini file (" + path + ").");
    return n;
  }
  double get_real(const std::string &section, const std::string &name) const {
    const std::string &v = get(section, name);
    char *end = nullptr;
    const double x = std::strtod(v.c_str(), &end);
    if (end == v.c_str())
      throw invalid_argument("The requested key (" + lower(section + "=" + name) +
                             ") does not seem to be a float in the .ini file (" + path + ").");
    return x;
  }
};

template <class T>
std::vector<T> cast_list(const std::vector<double> &v) {
  std::vector<T> out;
  out.reserve(v.size());
  for (double x : v) out.push_back(static_cast<T>(x));
  return out;
}

}  // namespace

std::map<std::string, std::string> read_ini_file(const std::string &path) {
  std::ifstream file(path);
  if (!file.good()) throw invalid_argument("Can't load .ini file.");  // src/fdModel.cpp:476-479
  std::map<std::string, std::string> values;
  std::string section, prev_name, raw;
  bool first = true;
  auto add = [&](const std::string &name, const std::string &value) {
    std::string &slot = values[lower(section + "=" + name)];
    if (!slot.empty()) slot += "\n";
    slot += value;
  };
  while (std::getline(file, raw)) {
    if (first) {  // UTF-8 BOM
      first = false;
      if (raw.size() >= 3 && static_cast<unsigned char>(raw[0]) == 0xEF &&
          static_cast<unsigned char>(raw[1]) == 0xBB && static_cast<unsigned char>(raw[2]) == 0xBF)
        raw = raw.substr(3);
    }
    const std::string line = rstrip(raw);
    const std::string start = lstrip(line);
    if (start.empty()) continue;
    if (start[0] == ';' || start[0] == '#') continue;
    if (!prev_name.empty() && start.size() < line.size()) {
      // indented non-blank line: continuation of the previous value
      add(prev_name, rstrip(cut_inline_comment(start)));
    } else if (start[0] == '[') {
      const size_t close = start.find(']');
      if (close == std::string::npos) break;  // the reference stops at the first syntax error
      section = start.substr(1, close - 1);
      prev_name.clear();
    } else {
      const size_t sep = start.find_first_of("=:");
      if (sep == std::string::npos) break;
      const std::string name = rstrip(start.substr(0, sep));
      const std::string value = rstrip(cut_inline_comment(lstrip(start.substr(sep + 1))));
      prev_name = name;
      add(name, value);
    }
  }
  return values;
}

std::vector<double> parse_list(std::string text) {
  strip_all_whitespace(text);
  const size_t close = text.find('}');
  if (close == std::string::npos || text.empty())
    throw invalid_argument("List value is not of the form {a, b, ...}: '" + text + "'");
  text.erase(close);
  text.erase(0, 1);  // leading brace
  std::vector<double> out;
  for (const std::string &tok : split(text, ",")) out.push_back(std::atof(tok.c_str()));
  return out;
}

std::vector<std::vector<int>> parse_nested_int_list(std::string text) {
  strip_all_whitespace(text);
  if (text.size() < 2) throw invalid_argument("Nested list value is not of the form {{a, b}, {c}}");
  text.erase(0, 2);
  std::vector<std::string> groups = split(text, "},{");
  std::string &last = groups.back();
  const size_t end = last.find("}};");
  if (end != std::string::npos) last.erase(end);
  std::vector<std::vector<int>> out;
  for (const std::string &g : groups) {
    std::vector<int> sub;
    for (const std::string &tok : split(g, ",")) sub.push_back(static_cast<int>(std::atof(tok.c_str())));
    out.push_back(sub);
  }
  return out;
}

void Config::finalize() {
  std::cout << "Parsing passed configuration." << std::endl;  // src/fdModel.cpp:396

  if (nt <= 0 || snapshot_interval <= 0 || nx_inner <= 0 || nz_inner <= 0 || np_boundary < 0 ||
      n_sources < 0 || n_shots < 0 || nr < 0 || npx <= 0 || npz <= 0)
    throw invalid_argument("Domain sizes, counts and intervals must be positive.");
  nx = nx_inner + 2 * np_boundary;
  nz = nz_inner + 2 * np_boundary;
  nx_free = nx_inner - 2 * nx_inner_boundary;
  nz_free = nz_inner - 2 * nz_inner_boundary;
  if (nx_free <= 0 || nz_free <= 0 || nx_inner_boundary < 0 || nz_inner_boundary < 0)
    throw invalid_argument("The free-parameter window is empty.");
  if (nx < 5 || nz < 5) throw invalid_argument("Grid too small for the 4th-order stencil.");
  // asserted (and compiled out) in the reference, src/fdModel.cpp:404-405
  if (nx_free % npx != 0 || nz_free % npz != 0)
    throw invalid_argument("basis.npx / basis.npz must divide the free-parameter window.");
  free_parameters = 3 * nx_free * nz_free / (npx * npz);

  if (static_cast<int>(ix_sources.size()) != n_sources ||
      static_cast<int>(iz_sources.size()) != n_sources ||
      static_cast<int>(moment_angles.size()) != n_sources)
    throw invalid_argument(
        "Dimension mismatch between n_sources and sources.ix_sources, "
        "sources.iz_sources or sources.moment_angles");
  if (static_cast<int>(which_source_to_fire_in_which_shot.size()) != n_shots)
    throw invalid_argument(
        "Mismatch between n_shots and sources.which_source_to_fire_in_which_shot");
  size_t total = 0;
  for (const auto &s : which_source_to_fire_in_which_shot) {
    total += s.size();
    for (int src : s)
      if (src < 0 || src >= n_sources)
        throw invalid_argument("sources.which_source_to_fire_in_which_shot names a source that does not exist.");
  }
  if (static_cast<int>(total) != n_sources)
    throw invalid_argument(
        "Mismatch between n_sources and sources.which_source_to_fire_in_which_shot.");
  if (static_cast<int>(ix_receivers.size()) != nr || static_cast<int>(iz_receivers.size()) != nr)
    throw invalid_argument(
        "Mismatch between nr and receivers.ix_receivers or receivers.iz_receivers");

  alpha = 1.0 / peak_frequency;
  snapshots = nt / snapshot_interval;  // integer division, src/fdModel.cpp:461
  if (std::floor(static_cast<double>(nt) / snapshot_interval) != snapshots)
    throw length_error("Snapshot interval and size of accumulator don't match!");

  if (!np_added) {  // src/fdModel.cpp:197-212
    for (int &v : ix_sources) v += np_boundary;
    for (int &v : iz_sources) v += np_boundary;
    for (int &v : ix_receivers) v += np_boundary;
    for (int &v : iz_receivers) v += np_boundary;
    np_added = true;
  }
  // The reference indexes blindly; out-of-grid stations would corrupt memory there.
  for (int s = 0; s < n_sources; ++s)
    if (ix_sources[s] < 1 || ix_sources[s] > nx - 2 || iz_sources[s] < 1 || iz_sources[s] > nz - 2)
      throw invalid_argument("A source (with its one-cell injection footprint) lies outside the grid.");
  for (int r = 0; r < nr; ++r)
    if (ix_receivers[r] < 0 || ix_receivers[r] >= nx || iz_receivers[r] < 0 || iz_receivers[r] >= nz)
      throw invalid_argument("A receiver lies outside the grid.");
}

Config config_from_ini(const std::string &path) {
  std::cout << "Loading configuration file: '" << path << "'." << std::endl;  // :472-473
  IniValues ini;
  ini.path = path;
  ini.values = read_ini_file(path);
  Config c;
  // same order of reads as src/fdModel.cpp:482-518, so the first missing key reported matches
  c.nt = static_cast<int>(ini.get_integer("domain", "nt"));
  c.nx_inner = static_cast<int>(ini.get_integer("domain", "nx_inner"));
  c.nz_inner = static_cast<int>(ini.get_integer("domain", "nz_inner"));
  c.nx_inner_boundary = static_cast<int>(ini.get_integer("domain", "nx_inner_boundary"));
  c.nz_inner_boundary = static_cast<int>(ini.get_integer("domain", "nz_inner_boundary"));
  c.dx = ini.get_real("domain", "dx");
  c.dz = ini.get_real("domain", "dz");
  c.dt = ini.get_real("domain", "dt");
  c.np_boundary = static_cast<int>(ini.get_integer("boundary", "np_boundary"));
  c.np_factor = ini.get_real("boundary", "np_factor");
  c.scalar_rho = ini.get_real("medium", "scalar_rho");
  c.scalar_vp = ini.get_real("medium", "scalar_vp");
  c.scalar_vs = ini.get_real("medium", "scalar_vs");
  c.npx = static_cast<int>(ini.get_integer("basis", "npx"));
  c.npz = static_cast<int>(ini.get_integer("basis", "npz"));
  c.peak_frequency = ini.get_real("sources", "peak_frequency");
  c.source_timeshift = ini.get_real("sources", "source_timeshift");
  c.delay_cycles_per_shot = ini.get_real("sources", "delay_cycles_per_shot");
  c.n_sources = static_cast<int>(ini.get_integer("sources", "n_sources"));
  c.n_shots = static_cast<int>(ini.get_integer("sources", "n_shots"));
  c.ix_sources = cast_list<int>(parse_list(ini.get("sources", "ix_sources")));
  c.iz_sources = cast_list<int>(parse_list(ini.get("sources", "iz_sources")));
  c.moment_angles = parse_list(ini.get("sources", "moment_angles"));
  c.which_source_to_fire_in_which_shot =
      parse_nested_int_list(ini.get("sources", "which_source_to_fire_in_which_shot"));
  c.nr = static_cast<int>(ini.get_integer("receivers", "nr"));
  c.ix_receivers = cast_list<int>(parse_list(ini.get("receivers", "ix_receivers")));
  c.iz_receivers = cast_list<int>(parse_list(ini.get("receivers", "iz_receivers")));
  c.snapshot_interval = static_cast<int>(ini.get_integer("inversion", "snapshot_interval"));
  c.observed_data_folder = ini.get("output", "observed_data_folder");
  c.stf_folder = ini.get("output", "stf_folder");
  c.finalize();
  return c;
}

}  // namespace psv
