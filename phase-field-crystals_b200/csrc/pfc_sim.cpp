// pfc_sim.cpp — native driver: the reference's executable phase_field_crystal_simulation_mpi
// (app/simulations/phase_field_crystal_simulation_mpi.cpp:15-183) and PhaseFieldCrystalSystemMPI<2>::run()
// (phase_field_crystal_system_mpi.cpp:650-711) on top of the C ABI, one process = one GPU = one rank.
//   pfc_sim <file.toml>            run the simulation (same keys, same pcout lines, same output file names)
//   pfc_sim --check <file.toml>    parse only: print what would be run (no GPU needed)
// deal.II's Triangulation / DoFHandler products come from the host stand-in (pfc_standin_*), everything numerical is one
// C-ABI call.  The parameter reader covers the value kinds of the reference's files: numbers, strings, booleans,
// arrays of numbers and arrays of arrays of numbers, `#` comments, an optional [gpu] table with pfc_params fields.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include <sys/stat.h>

#include "../../include/pfc_b200.h"

namespace
{
struct Value
{
  enum Kind { NUMBER, STRING, BOOL, ARRAY } kind = NUMBER;
  double number = 0;
  bool is_integer = false;
  std::string string;
  bool boolean = false;
  std::vector<Value> array;
};
using Table = std::map<std::string, Value>;

struct Parser
{
  const std::string &s;
  size_t i = 0;
  int line = 1;
  explicit Parser(const std::string &text) : s(text) {}
  [[noreturn]] void fail(const std::string &what) const
  {
    throw std::runtime_error("Parsing failed:\nline " + std::to_string(line) + ": " + what);
  }
  void skip_space(bool newlines)
  {
    while (i < s.size())
      {
        if (s[i] == '#')
          while (i < s.size() && s[i] != '\n')
            ++i;
        else if (s[i] == ' ' || s[i] == '\t' || s[i] == '\r')
          ++i;
        else if (s[i] == '\n' && newlines)
          {
            ++line;
            ++i;
          }
        else
          break;
      }
  }
  Value value()
  {
    skip_space(false);
    Value v;
    if (i >= s.size())
      fail("value expected");
    if (s[i] == '"')
      {
        v.kind = Value::STRING;
        for (++i; i < s.size() && s[i] != '"'; ++i)
          {
            if (s[i] == '\n')
              fail("unterminated string");
            v.string += s[i];
          }
        if (i >= s.size())
          fail("unterminated string");
        ++i;
      }
    else if (s[i] == '[')
      {
        v.kind = Value::ARRAY;
        ++i;
        for (;;)
          {
            skip_space(true);
            if (i < s.size() && s[i] == ']')
              {
                ++i;
                break;
              }
            v.array.push_back(value());
            skip_space(true);
            if (i < s.size() && s[i] == ',')
              ++i;
            else if (i < s.size() && s[i] == ']')
              {
                ++i;
                break;
              }
            else
              fail("',' or ']' expected");
          }
      }
    else if (s.compare(i, 4, "true") == 0)
      {
        v.kind = Value::BOOL;
        v.boolean = true;
        i += 4;
      }
    else if (s.compare(i, 5, "false") == 0)
      {
        v.kind = Value::BOOL;
        i += 5;
      }
    else
      {
        size_t j = i;
        while (j < s.size() && (std::isdigit((unsigned char)s[j]) || std::strchr("+-.eE_", s[j])))
          ++j;
        std::string tok = s.substr(i, j - i);
        if (tok.empty())
          fail("value expected");
        v.is_integer = tok.find_first_of(".eE") == std::string::npos;
        std::string clean;
        for (char c : tok)
          if (c != '_')
            clean += c;
        char *end = nullptr;
        v.number = std::strtod(clean.c_str(), &end);
        if (!end || *end)
          fail("bad number '" + tok + "'");
        i = j;
      }
    return v;
  }
  // root table + sub tables as "table.key"
  Table parse()
  {
    Table t;
    std::string prefix;
    for (;;)
      {
        skip_space(true);
        if (i >= s.size())
          break;
        if (s[i] == '[')
          {
            const size_t e = s.find(']', i);
            if (e == std::string::npos)
              fail("']' expected");
            prefix = s.substr(i + 1, e - i - 1) + ".";
            i = e + 1;
            continue;
          }
        size_t j = i;
        while (j < s.size() && (std::isalnum((unsigned char)s[j]) || s[j] == '_' || s[j] == '-'))
          ++j;
        if (j == i)
          fail("key expected");
        const std::string key = prefix + s.substr(i, j - i);
        i = j;
        skip_space(false);
        if (i >= s.size() || s[i] != '=')
          fail("'=' expected after " + key);
        ++i;
        t[key] = value();
        skip_space(false);
        if (i < s.size() && s[i] != '\n')
          fail("end of line expected after the value of " + key);
      }
    return t;
  }
};

struct Simulation // the 15 constructor arguments of mpi.hpp:30-50 + HexagonalLattice's
{
  unsigned degree = 0, output_interval = 0, n_timesteps = 0, simulation_max_iters = 0, n_refines = 0;
  std::string data_folder, configuration_stem, rhs_stem;
  double eps = 0, dt = 0, theta = 0, simulation_tol = 0, p1[2] = {0, 0}, p2[2] = {0, 0}, psi_0 = 0, A_0 = 0;
  std::vector<double> dislocation_positions, burgers_vectors; // [n][2]
  pfc_params gpu;
};

std::string stem(const std::string &filename)
{
  const size_t slash = filename.find_last_of('/');
  std::string base = slash == std::string::npos ? filename : filename.substr(slash + 1);
  const size_t dot = base.find_last_of('.');
  return (dot == std::string::npos || dot == 0) ? base : base.substr(0, dot);
}

const Value *find(const Table &t, const char *key, Value::Kind kind)
{
  const auto it = t.find(key);
  return (it != t.end() && it->second.kind == kind) ? &it->second : nullptr;
}
double number(const Table &t, const char *key, const char *what)
{
  const Value *v = find(t, key, Value::NUMBER);
  if (!v)
    throw std::invalid_argument(std::string("No ") + what + " in parameter file"); // app_mpi.cpp:25-28, 45-52
  return v->number;
}
std::string text(const Table &t, const char *key, const char *what)
{
  const Value *v = find(t, key, Value::STRING);
  if (!v)
    throw std::invalid_argument(std::string("No ") + what + " in parameter file");
  return v->string;
}
std::vector<double> points(const Table &t, const char *key, const char *message, double scale)
{
  const Value *v = find(t, key, Value::ARRAY);
  if (!v)
    throw std::invalid_argument(message);
  std::vector<double> out;
  for (const Value &p : v->array)
    {
      if (p.kind != Value::ARRAY || p.array.size() != 2 || p.array[0].kind != Value::NUMBER || p.array[1].kind != Value::NUMBER)
        throw std::invalid_argument(message);
      out.push_back(p.array[0].number * scale);
      out.push_back(p.array[1].number * scale);
    }
  return out;
}

// parse_simulation_parameters<2> (app_mpi.cpp:15-141)
Simulation parse_simulation_parameters(const Table &t)
{
  Simulation s;
  s.data_folder = text(t, "data_folder", "data folder");
  s.configuration_stem = stem(text(t, "configuration_filename", "configuration filename"));
  s.rhs_stem = stem(text(t, "rhs_filename", "rhs filename"));
  s.output_interval = (unsigned)number(t, "output_interval", "output_interval");
  s.degree = (unsigned)number(t, "degree", "degree");
  s.eps = number(t, "eps", "eps");
  s.dt = number(t, "dt", "dt");
  s.n_timesteps = (unsigned)number(t, "n_timesteps", "n_timesteps");
  s.theta = number(t, "theta", "theta");
  s.simulation_tol = number(t, "simulation_tol", "simulation_tol");
  s.simulation_max_iters = (unsigned)number(t, "simulation_max_iters", "simulation_max_iters");
  s.n_refines = (unsigned)number(t, "n_refines", "n_refines");
  const Value *sc = find(t, "scale_by_lattice_constant", Value::BOOL);
  const double a = 4.0 * M_PI / std::sqrt(3.0); // HexagonalLattice<dim>::a (hexagonal_lattice.hpp:35)
  const double scale = (sc && sc->boolean) ? a : 1.0;
  for (int k = 0; k < 2; ++k)
    {
      const Value *p = find(t, k == 0 ? "p1" : "p2", Value::ARRAY);
      if (!p || p->array.size() != 2 || p->array[0].kind != Value::NUMBER || p->array[1].kind != Value::NUMBER)
        throw std::invalid_argument("p1 is not an array!"); // the reference's text for both, app_mpi.cpp:61,69
      (k == 0 ? s.p1 : s.p2)[0] = p->array[0].number * scale;
      (k == 0 ? s.p1 : s.p2)[1] = p->array[1].number * scale;
    }
  s.psi_0 = number(t, "psi_0", "psi_0");
  if (find(t, "A_0", Value::NUMBER))
    s.A_0 = find(t, "A_0", Value::NUMBER)->number;
  else if (find(t, "A_0", Value::STRING) && find(t, "A_0", Value::STRING)->string == "default")
    s.A_0 = 0.2 * (std::abs(s.psi_0) + (1.0 / 3.0) * std::sqrt(-15 * s.eps - 36 * s.psi_0 * s.psi_0)); // app_mpi.cpp:87-88
  else
    throw std::invalid_argument("Incorrect input for A_0");
  s.dislocation_positions = points(t, "dislocation_positions", "Incorrect input for dislocation positions", scale);
  s.burgers_vectors = points(t, "burgers_vectors", "Incorrect input for burgers vectors", scale);
  if (s.dislocation_positions.size() != s.burgers_vectors.size())
    throw std::invalid_argument("dislocation_positions and burgers_vectors differ in length");

  pfc_default_params(&s.gpu);
  s.gpu.degree = (int32_t)s.degree;
  s.gpu.dt = s.dt;
  s.gpu.theta = s.theta;
  s.gpu.eps = s.eps;
  s.gpu.newton_tol = s.simulation_tol;
  s.gpu.newton_max_iters = (int32_t)s.simulation_max_iters;
  s.gpu.operator_kind = PFC_OPERATOR_MATRIX_FREE; // lattice path + block-Schur preconditioner unless [gpu] says otherwise
  s.gpu.preconditioner = PFC_PRECOND_BLOCK_SCHUR;
  struct
  {
    const char *key;
    int32_t *field;
  } opts[] = {{"gpu.operator_kind", &s.gpu.operator_kind}, {"gpu.preconditioner", &s.gpu.preconditioner},
              {"gpu.precond_sweeps", &s.gpu.precond_sweeps}, {"gpu.orthogonalization", &s.gpu.orthogonalization},
              {"gpu.use_cuda_graph", &s.gpu.use_cuda_graph}, {"gpu.device", &s.gpu.device}};
  for (auto &o : opts)
    if (const Value *v = find(t, o.key, Value::NUMBER))
      *o.field = (int32_t)v->number;
  return s;
}

void check(int rc, pfc_handle *h, const char *what)
{
  if (rc != PFC_OK && rc != PFC_WARN_NEWTON_MAX_ITERS) // Newton hitting its cap is accepted silently (mpi.cpp:579-592)
    throw std::runtime_error(std::string(what) + ": " + pfc_last_error(h));
}

// PhaseFieldCrystalSystemMPI<2>::run (mpi.cpp:650-711)
void run(const Simulation &s)
{
  pfc_handle *h = nullptr;
  check(pfc_create(&s.gpu, &h), nullptr, "pfc_create");
  if (!s.data_folder.empty())
    mkdir(s.data_folder.c_str(), 0777); // (deal.II would fail on a missing folder)
  pfc_standin *grid = nullptr;
  try
    {
      std::printf("Making grid!\n");
      if (pfc_standin_create((int32_t)s.degree, (int32_t)s.n_refines, s.p1[0], s.p1[1], s.p2[0], s.p2[1], 0, 1, &grid) != PFC_OK)
        throw std::runtime_error("pfc_standin_create failed");
      const pfc_mesh *mesh = pfc_standin_mesh(grid);
      std::printf("Setting up dofs!\n");
      check(pfc_setup(h, mesh), h, "pfc_setup");
      std::printf("Initializing fe field!\n");
      const int64_t N = pfc_standin_n_nodes(grid);
      std::vector<double> Psi((size_t)3 * N);
      if (pfc_standin_hexagonal_lattice(grid, s.A_0, s.psi_0, (int32_t)(s.dislocation_positions.size() / 2),
                                        s.dislocation_positions.data(), s.burgers_vectors.data(), Psi.data()) != PFC_OK)
        throw std::runtime_error("pfc_standin_hexagonal_lattice failed");
      check(pfc_set_state(h, Psi.data(), Psi.data()), h, "pfc_set_state");

      const char *dir = s.data_folder.c_str();
      std::printf("Outputting initial configuration!\n");
      check(pfc_output(h, mesh, 0, dir, s.configuration_stem.c_str(), 0), h, "pfc_output");
      std::printf("Outputting initial right-hand side!\n");
      {
        // system_rhs is all zero before the first assembly (mpi.cpp:271)
        std::vector<double> zeros((size_t)3 * N, 0.0);
        const char *names[3] = {"psi", "chi", "phi"};
        char err[400] = {0};
        if (pfc_vtu_write(mesh, dir, s.rhs_stem.c_str(), 0, 3, names, nullptr, zeros.data(), err, sizeof(err)) != PFC_OK)
          throw std::runtime_error(err);
      }
      std::printf("Initializing stress_calculator!\n");
      std::vector<double> stress((size_t)4 * N);
      int64_t its[4];
      bool have_stress = true;
      auto stress_step = [&](bool announce) {
        if (!have_stress)
          return;
        const int rc = pfc_stress_owned(h, s.eps, 1e-10, stress.data(), its);
        if (rc == PFC_ERR_UNSUPPORTED)
          {
            have_stress = false;
            std::printf("stress projection needs the lattice path (operator_kind = 1): skipped\n");
            return;
          }
        check(rc, h, "pfc_stress");
        if (announce)
          for (int c = 0; c < 4; ++c)
            std::printf("Number of stress iterations: %lld\n", (long long)its[c]); // mpi.cpp:677-678
        check(pfc_output(h, mesh, 2, dir, "stress", 0), h, "pfc_output");
      };
      stress_step(true);

      double ms_assembly = 0, ms_solver = 0, ms_total = 0;
      for (unsigned timestep = 1; timestep <= s.n_timesteps; ++timestep)
        {
          std::printf("Iterating timestep: %u\n", timestep);
          pfc_step_stats st;
          const int rc = pfc_step(h, &st);
          for (int i = 0; i < st.n_assemblies; ++i)
            {
              std::printf("Residual is: %.17g\n", st.residual[i]); // mpi.cpp:585
              if (i < st.n_solves)
                std::printf("Number of iterations: %lld\n", (long long)st.gmres_iters[i]); // mpi.cpp:483
            }
          check(rc, h, "pfc_step");
          ms_assembly += st.ms_assembly;
          ms_solver += st.ms_solver;
          ms_total += st.ms_total;
          stress_step(false);
          if (timestep % s.output_interval == 0)
            {
              std::printf("Outputting configuration!\n");
              check(pfc_output(h, mesh, 0, dir, s.configuration_stem.c_str(), timestep), h, "pfc_output");
              std::printf("Outputting right-hand side!\n");
              check(pfc_output(h, mesh, 1, dir, s.rhs_stem.c_str(), timestep), h, "pfc_output");
            }
          std::printf("\n");
        }
      // timer.print_summary() (mpi.cpp:711): the sections the hot path owns
      std::printf("assembly %.3f ms | solver %.3f ms | timesteps total %.3f ms\n", ms_assembly, ms_solver, ms_total);
    }
  catch (...)
    {
      if (grid)
        pfc_standin_destroy(grid);
      pfc_destroy(h);
      throw;
    }
  pfc_standin_destroy(grid);
  pfc_destroy(h);
}
} // namespace

int main(int ac, char *av[]) // app_mpi.cpp:143-183: errors are printed, the exit code stays 0; a parse error returns 1
{
  try
    {
      const bool check_only = ac > 2 && std::strcmp(av[1], "--check") == 0;
      const char *file = ac > 1 ? av[check_only ? 2 : 1] : nullptr;
      if (!file)
        throw std::invalid_argument("usage: pfc_sim [--check] <file.toml>");
      std::ifstream in(file);
      if (!in)
        throw std::invalid_argument(std::string("cannot open ") + file);
      std::stringstream buf;
      buf << in.rdbuf();
      const std::string content = buf.str();
      Table tbl;
      try
        {
          tbl = Parser(content).parse();
        }
      catch (const std::runtime_error &err)
        {
          std::fprintf(stderr, "%s\n", err.what());
          return 1;
        }
      const Value *dim = find(tbl, "dim", Value::NUMBER);
      if (dim && dim->number == 2)
        {
          const Simulation s = parse_simulation_parameters(tbl);
          if (check_only)
            {
              std::printf("degree %u n_refines %u n_timesteps %u output_interval %u\n", s.degree, s.n_refines, s.n_timesteps, s.output_interval);
              std::printf("eps %.17g dt %.17g theta %.17g simulation_tol %.17g simulation_max_iters %u\n", s.eps, s.dt, s.theta,
                          s.simulation_tol, s.simulation_max_iters);
              std::printf("p1 %.17g %.17g p2 %.17g %.17g\n", s.p1[0], s.p1[1], s.p2[0], s.p2[1]);
              std::printf("psi_0 %.17g A_0 %.17g dislocations %zu\n", s.psi_0, s.A_0, s.dislocation_positions.size() / 2);
              std::printf("files %s%s_NN.R.vtu %s%s_NN.R.vtu\n", s.data_folder.c_str(), s.configuration_stem.c_str(),
                          s.data_folder.c_str(), s.rhs_stem.c_str());
              std::printf("gpu operator_kind %d preconditioner %d\n", s.gpu.operator_kind, s.gpu.preconditioner);
            }
          else
            run(s);
        }
      else if (dim && dim->number == 3)
        throw std::invalid_argument("dim = 3 is instantiated upstream (mpi.cpp:716) but not supported by this path");
    }
  catch (const std::exception &err)
    {
      std::fprintf(stderr, "--------------------------\nFound error in program!\n%s--------------------------\n", err.what());
    }
  return 0;
}
