// pfc_vtu.cpp — host-side VTU / PVTU writer.
// Replaces what the reference gets from deal.II's DataOut in output_configuration / output_rhs
// (phase_field_crystal_system_mpi.cpp:597-645) and StressCalculatorMPI::output_stress
// (stress_calculator_mpi.cpp:358-384):
//   data_out.build_patches()             -> one VTK_QUAD per locally owned cell carrying its 4 vertex values
//                                           (default n_subdivisions = 1: Q2 edge/interior nodes are not written)
//   write_vtu_with_pvtu_record(dir, base, counter, comm, 2)
//                                        -> <dir><base>_<counter, 2 digits>.<rank>.vtu per rank and
//                                           <dir><base>_<counter, 2 digits>.pvtu on rank 0
// Fields keep the reference's names ("psi", "chi", "phi"; tensor "stress" as a 9-component array, the 2x2 block
// filled, as DataOut pads component_is_part_of_tensor data).  Arrays are raw appended binary (Float64, little
// endian, UInt64 block headers); deal.II itself writes zlib-compressed base64 — same information, same names.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include "../../include/pfc_b200.h"

namespace
{
std::string int_to_string(unsigned v, unsigned digits)
{
  char buf[32];
  std::snprintf(buf, sizeof(buf), "%0*u", (int)digits, v);
  return buf;
}
unsigned needed_digits(unsigned max_number)
{
  unsigned d = 1;
  while (max_number >= 10)
    {
      max_number /= 10;
      ++d;
    }
  return d;
}
int set_err(char *err, int len, const std::string &m)
{
  if (err && len > 0)
    {
      std::strncpy(err, m.c_str(), (size_t)len - 1);
      err[len - 1] = 0;
    }
  return PFC_ERR_BAD_ARGUMENT;
}

struct Group // one output array: a scalar, or the fields of one tensor
{
  std::string name;
  std::vector<int> fields;
  bool tensor;
};
std::vector<Group> group_fields(int n_fields, const char *const *names, const int32_t *interp)
{
  std::vector<Group> g;
  for (int f = 0; f < n_fields; ++f)
    {
      const bool t = interp && interp[f] == 1;
      if (t && !g.empty() && g.back().tensor && g.back().name == names[f] && g.back().fields.size() < 4)
        g.back().fields.push_back(f);
      else
        g.push_back({names[f], {f}, t});
    }
  return g;
}
} // namespace

extern "C" int pfc_vtu_piece_name(const char *basename, uint32_t counter, int32_t rank, int32_t n_ranks, char *out, int32_t outlen)
{
  if (!basename || !out || outlen <= 0 || n_ranks < 1)
    return PFC_ERR_BAD_ARGUMENT;
  const std::string s = std::string(basename) + "_" + int_to_string(counter, 2) +
                        (rank < 0 ? std::string(".pvtu")
                                  : "." + int_to_string((unsigned)rank, needed_digits((unsigned)std::max(0, n_ranks - 1))) + ".vtu");
  if ((int)s.size() + 1 > outlen)
    return PFC_ERR_BAD_ARGUMENT;
  std::memcpy(out, s.c_str(), s.size() + 1);
  return PFC_OK;
}

extern "C" int pfc_vtu_write(const pfc_mesh *mesh, const char *directory, const char *basename, uint32_t counter,
                             int32_t n_fields, const char *const *names, const int32_t *interpretation,
                             const double *values, char *err, int32_t errlen)
{
  if (!mesh || !directory || !basename || n_fields < 1 || !names || !values)
    return set_err(err, errlen, "pfc_vtu_write: null argument");
  const int dpc = mesh->dofs_per_cell, n_s = dpc / 3;
  const int64_t N = mesh->block_size[0], nc = mesh->n_owned_cells;
  // local system index of the psi dof on vertex v (deal.II FE_Q numbering: the first four scalar shapes are the
  // vertices in lexicographic order)
  int vert[4] = {-1, -1, -1, -1};
  for (int i = 0; i < dpc; ++i)
    if (mesh->local_component[i] == 0 && mesh->local_scalar_index[i] < 4)
      vert[mesh->local_scalar_index[i]] = i;
  if (n_s < 4 || vert[0] < 0 || vert[1] < 0 || vert[2] < 0 || vert[3] < 0)
    return set_err(err, errlen, "pfc_vtu_write: no vertex dofs in the local numbering");
  const std::vector<Group> groups = group_fields(n_fields, names, interpretation);

  char piece[512], record[512];
  if (pfc_vtu_piece_name(basename, counter, mesh->rank, mesh->n_ranks, piece, sizeof(piece)) != PFC_OK ||
      pfc_vtu_piece_name(basename, counter, -1, mesh->n_ranks, record, sizeof(record)) != PFC_OK)
    return set_err(err, errlen, "pfc_vtu_write: file name too long");
  // like the reference, directory and file name are concatenated as they are (mpi.cpp:612-616)
  const std::string path = std::string(directory) + piece;
  std::ofstream out(path, std::ios::binary);
  if (!out)
    return set_err(err, errlen, "pfc_vtu_write: cannot open " + path);

  const uint64_t np = 4 * (uint64_t)nc;
  uint64_t offset = 0;
  auto block = [&](uint64_t bytes) {
    const uint64_t o = offset;
    offset += 8 + bytes;
    return o;
  };
  out << "<?xml version=\"1.0\"?>\n<VTKFile type=\"UnstructuredGrid\" version=\"1.0\" byte_order=\"LittleEndian\" header_type=\"UInt64\">\n"
      << "<UnstructuredGrid>\n<Piece NumberOfPoints=\"" << np << "\" NumberOfCells=\"" << nc << "\">\n";
  out << "<Points>\n<DataArray type=\"Float64\" NumberOfComponents=\"3\" format=\"appended\" offset=\"" << block(np * 24) << "\"/>\n</Points>\n";
  out << "<Cells>\n<DataArray type=\"Int64\" Name=\"connectivity\" format=\"appended\" offset=\"" << block(np * 8) << "\"/>\n";
  out << "<DataArray type=\"Int64\" Name=\"offsets\" format=\"appended\" offset=\"" << block((uint64_t)nc * 8) << "\"/>\n";
  out << "<DataArray type=\"UInt8\" Name=\"types\" format=\"appended\" offset=\"" << block((uint64_t)nc) << "\"/>\n</Cells>\n";
  out << "<PointData Scalars=\"scalars\">\n";
  for (const Group &g : groups)
    out << "<DataArray type=\"Float64\" Name=\"" << g.name << "\"" << (g.tensor ? " NumberOfComponents=\"9\"" : "")
        << " format=\"appended\" offset=\"" << block(np * 8 * (g.tensor ? 9 : 1)) << "\"/>\n";
  out << "</PointData>\n</Piece>\n</UnstructuredGrid>\n<AppendedData encoding=\"raw\">\n_";

  auto header = [&](uint64_t bytes) { out.write(reinterpret_cast<const char *>(&bytes), 8); };
  {
    std::vector<double> pts(np * 3);
    for (int64_t k = 0; k < nc; ++k)
      for (int v = 0; v < 4; ++v)
        {
          pts[(4 * k + v) * 3 + 0] = mesh->cell_vertices[(k * 4 + v) * 2 + 0];
          pts[(4 * k + v) * 3 + 1] = mesh->cell_vertices[(k * 4 + v) * 2 + 1];
          pts[(4 * k + v) * 3 + 2] = 0.0;
        }
    header(np * 24);
    out.write(reinterpret_cast<const char *>(pts.data()), (std::streamsize)(np * 24));
  }
  {
    std::vector<int64_t> conn(np), offs((size_t)nc);
    static const int quad[4] = {0, 1, 3, 2}; // VTK_QUAD runs counter-clockwise, the patch is lexicographic
    for (int64_t k = 0; k < nc; ++k)
      {
        for (int v = 0; v < 4; ++v)
          conn[4 * k + v] = 4 * k + quad[v];
        offs[k] = 4 * (k + 1);
      }
    header(np * 8);
    out.write(reinterpret_cast<const char *>(conn.data()), (std::streamsize)(np * 8));
    header((uint64_t)nc * 8);
    out.write(reinterpret_cast<const char *>(offs.data()), (std::streamsize)(nc * 8));
    std::vector<uint8_t> types((size_t)nc, 9);
    header((uint64_t)nc);
    out.write(reinterpret_cast<const char *>(types.data()), (std::streamsize)nc);
  }
  for (const Group &g : groups)
    {
      const int ncomp = g.tensor ? 9 : 1;
      std::vector<double> data(np * ncomp, 0.0);
      for (int64_t k = 0; k < nc; ++k)
        for (int v = 0; v < 4; ++v)
          {
            const int64_t t = mesh->cell_dofs[k * dpc + vert[v]] - mesh->block_start[0];
            if (t < 0 || t >= N)
              return set_err(err, errlen, "pfc_vtu_write: vertex dof outside the psi block");
            if (!g.tensor)
              data[4 * k + v] = values[(int64_t)g.fields[0] * N + t];
            else
              for (size_t c = 0; c < g.fields.size(); ++c) // 2-D tensor component c = a*2 + b -> slot a*3 + b
                data[(4 * k + v) * 9 + (c / 2) * 3 + (c % 2)] = values[(int64_t)g.fields[c] * N + t];
          }
      header(np * 8 * ncomp);
      out.write(reinterpret_cast<const char *>(data.data()), (std::streamsize)(np * 8 * ncomp));
    }
  out << "\n</AppendedData>\n</VTKFile>\n";
  out.close();
  if (!out)
    return set_err(err, errlen, "pfc_vtu_write: write failed for " + path);

  if (mesh->rank == 0)
    {
      const std::string rpath = std::string(directory) + record;
      std::ofstream r(rpath);
      if (!r)
        return set_err(err, errlen, "pfc_vtu_write: cannot open " + rpath);
      r << "<?xml version=\"1.0\"?>\n<VTKFile type=\"PUnstructuredGrid\" version=\"1.0\" byte_order=\"LittleEndian\" header_type=\"UInt64\">\n"
        << "<PUnstructuredGrid GhostLevel=\"0\">\n<PPointData Scalars=\"scalars\">\n";
      for (const Group &g : groups)
        r << "<PDataArray type=\"Float64\" Name=\"" << g.name << "\"" << (g.tensor ? " NumberOfComponents=\"9\"" : "") << "/>\n";
      r << "</PPointData>\n<PPoints>\n<PDataArray type=\"Float64\" NumberOfComponents=\"3\"/>\n</PPoints>\n";
      for (int q = 0; q < mesh->n_ranks; ++q)
        {
          char pn[512];
          pfc_vtu_piece_name(basename, counter, q, mesh->n_ranks, pn, sizeof(pn));
          r << "<Piece Source=\"" << pn << "\"/>\n";
        }
      r << "</PUnstructuredGrid>\n</VTKFile>\n";
      if (!r)
        return set_err(err, errlen, "pfc_vtu_write: write failed for " + rpath);
    }
  return PFC_OK;
}
