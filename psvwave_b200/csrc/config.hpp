// Host-side experiment description: .ini reading, validation and derived geometry.
// Replaces fdModel::parse_configuration_file / parse_parameters
// (reference src/fdModel.cpp:390-523) and the list helpers (:1574-1659).
#pragma once

#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace psv {

// The two exception types the reference throws on this path (mapped to status codes in
// capi.cu and back to Python ValueError by the binding).
using invalid_argument = std::invalid_argument;
using length_error = std::length_error;

struct Config {
  // [domain]
  int nt = 0, nx_inner = 0, nz_inner = 0, nx_inner_boundary = 0, nz_inner_boundary = 0;
  double dx = 0, dz = 0, dt = 0;
  // [boundary]
  int np_boundary = 0;
  double np_factor = 0;
  // [medium]
  double scalar_rho = 0, scalar_vp = 0, scalar_vs = 0;
  // [basis]
  int npx = 1, npz = 1;
  // [sources]
  double peak_frequency = 0, source_timeshift = 0, delay_cycles_per_shot = 0;
  int n_sources = 0, n_shots = 0;
  std::vector<int> ix_sources, iz_sources;  // physical-domain indices until finalize()
  std::vector<double> moment_angles;
  std::vector<std::vector<int>> which_source_to_fire_in_which_shot;
  // [receivers]
  int nr = 0;
  std::vector<int> ix_receivers, iz_receivers;
  // [inversion]
  int snapshot_interval = 1;
  // [output]
  std::string observed_data_folder = ".", stf_folder = ".";

  // derived (src/fdModel.cpp:398-407,460-461)
  int nx = 0, nz = 0, nx_free = 0, nz_free = 0, free_parameters = 0, snapshots = 0;
  double alpha = 0;
  bool np_added = false;  // source/receiver indices already include np_boundary

  // Validation + derived sizes; throws like parse_parameters (src/fdModel.cpp:414-467).
  // Adds np_boundary to the source/receiver indices once (src/fdModel.cpp:197-212).
  void finalize();
};

// .ini text -> lower-cased "section=name" -> value.  Same line rules as the inih fork the
// reference vendors (ext/inih/INIReader.h:186-283): ';' or '#' comment lines, inline ';'
// comments only after whitespace, name[=:]value, continuation lines (leading whitespace)
// and repeated keys joined with '\n', section/name lookups case-insensitive.  One
// deliberate superset: no 200-character line limit (ext/inih/INIReader.h:90-92), so long
// receiver lists work.
std::map<std::string, std::string> read_ini_file(const std::string &path);

// {a, b, c} lists (src/fdModel.cpp:1574-1599) and {{0,1},{2}} nested lists (:1601-1659).
std::vector<double> parse_list(std::string text);
std::vector<std::vector<int>> parse_nested_int_list(std::string text);

// Full reference behaviour of fdModel(const char*) up to and including parse_parameters.
// Prints the reference's two progress lines (src/fdModel.cpp:396,472-473).
Config config_from_ini(const std::string &path);

}  // namespace psv
