// pf_io.inl - the reference host's output files, written by plain host code (no device): maps_NNNN.bin
// (reference main.cpp:705-735, read by plot.py:26-73), stats.csv (main.cpp:741-757, 782-825, read by plot.py:104-202)
// and a minimal PNG writer for the headless view. Errors go to pf_last_error(NULL).
#include <errno.h>

namespace {

const int32_t PF_BIN_FILE_MAGIC = 0x11223344;      // BIN_FILE_MAGIC, main.cpp:703
enum { PF_MAP_NAME_BYTES = 32 };                   // sizeof Sim_Map::name, simulation.h:50

size_t maps_bin_header_bytes(int n_maps) { return 4 * 4 + 3 * 8 + 8 + (size_t) n_maps * PF_MAP_NAME_BYTES; }

int write_maps_header(FILE* f, int nx, int ny, double dx, double dy, int64_t iter, double time, int n_maps, const char* const* names)
{
    int32_t i4[4] = {PF_BIN_FILE_MAGIC, n_maps, nx, ny};
    double d3[3] = {dx, dy, time};
    fwrite(i4, sizeof i4, 1, f);
    fwrite(d3, sizeof d3, 1, f);
    fwrite(&iter, sizeof iter, 1, f);
    for(int i = 0; i < n_maps; i++)
    {
        char name[PF_MAP_NAME_BYTES] = {0};            // sim_realloc: strncpy + terminator, simulation.cu:97-98
        strncpy(name, names[i] ? names[i] : "", sizeof name - 1);
        fwrite(name, sizeof name, 1, f);
    }
    return ferror(f) ? -1 : 0;
}

// CRC-32 (PNG / zlib polynomial) and Adler-32
uint32_t crc32_update(uint32_t crc, const uint8_t* p, size_t n)
{
    static uint32_t table[256]; static bool init = false;
    if(!init) { for(uint32_t i = 0; i < 256; i++) { uint32_t c = i; for(int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1; table[i] = c; } init = true; }
    crc = ~crc;
    for(size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xFF] ^ (crc >> 8);
    return ~crc;
}

void put_be32(uint8_t* p, uint32_t v) { p[0] = (uint8_t)(v >> 24); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t) v; }

int png_chunk(FILE* f, const char type[4], const uint8_t* data, size_t n)
{
    uint8_t hdr[8]; put_be32(hdr, (uint32_t) n); memcpy(hdr + 4, type, 4);
    uint32_t crc = crc32_update(0, hdr + 4, 4);
    if(n) crc = crc32_update(crc, data, n);
    uint8_t tail[4]; put_be32(tail, crc);
    fwrite(hdr, 8, 1, f);
    if(n) fwrite(data, n, 1, f);
    fwrite(tail, 4, 1, f);
    return ferror(f) ? -1 : 0;
}

} // namespace

extern "C" {

int pf_write_maps_bin_header(const char* path, int nx, int ny, double dx, double dy, int64_t iter, double time, int n_maps, const char* const* names)
{
    if(!path || nx <= 0 || ny <= 0 || n_maps < 0 || (n_maps > 0 && !names)) return fail(nullptr, PF_ERR_INVALID, "pf_write_maps_bin_header: bad argument");
    FILE* f = fopen(path, "wb");
    if(!f) return fail(nullptr, PF_ERR_INVALID, "cannot open '%s': %s", path, strerror(errno));
    int rc = write_maps_header(f, nx, ny, dx, dy, iter, time, n_maps, names);
    // size the file for the maps so that rows can land in any order
    const long long total = (long long) maps_bin_header_bytes(n_maps) + (long long) n_maps * nx * ny * 8LL;
    if(rc == 0 && n_maps > 0 && (fseeko(f, (off_t)(total - 1), SEEK_SET) != 0 || fputc(0, f) == EOF)) rc = -1;
    if(fclose(f) != 0) rc = -1;
    return rc ? fail(nullptr, PF_ERR_INVALID, "error writing '%s'", path) : PF_OK;
}

int pf_write_maps_bin_rows(const char* path, int nx, int ny, int n_maps, int map_index, int y0, int rows, const double* data)
{
    if(!path || !data || nx <= 0 || ny <= 0 || map_index < 0 || map_index >= n_maps || y0 < 0 || rows < 0 || y0 + rows > ny)
        return fail(nullptr, PF_ERR_INVALID, "pf_write_maps_bin_rows: bad argument");
    FILE* f = fopen(path, "r+b");
    if(!f) return fail(nullptr, PF_ERR_INVALID, "cannot open '%s': %s", path, strerror(errno));
    const long long off = (long long) maps_bin_header_bytes(n_maps) + ((long long) map_index * ny + y0) * nx * 8LL;
    int rc = fseeko(f, (off_t) off, SEEK_SET) != 0 ? -1 : 0;
    if(rc == 0 && rows > 0 && fwrite(data, sizeof(double) * (size_t) nx, (size_t) rows, f) != (size_t) rows) rc = -1;
    if(fclose(f) != 0) rc = -1;
    return rc ? fail(nullptr, PF_ERR_INVALID, "error writing '%s'", path) : PF_OK;
}

int pf_write_maps_bin(const char* path, int nx, int ny, double dx, double dy, int64_t iter, double time, int n_maps, const char* const* names, const double* const* maps)
{
    if(!path || nx <= 0 || ny <= 0 || n_maps < 0 || (n_maps > 0 && (!names || !maps))) return fail(nullptr, PF_ERR_INVALID, "pf_write_maps_bin: bad argument");
    FILE* f = fopen(path, "wb");
    if(!f) return fail(nullptr, PF_ERR_INVALID, "cannot open '%s': %s", path, strerror(errno));
    int rc = write_maps_header(f, nx, ny, dx, dy, iter, time, n_maps, names);
    const size_t N = (size_t) nx * (size_t) ny;
    for(int i = 0; i < n_maps && rc == 0; i++)
        if(!maps[i] || fwrite(maps[i], sizeof(double), N, f) != N) rc = -1;
    if(fclose(f) != 0) rc = -1;
    return rc ? fail(nullptr, PF_ERR_INVALID, "error writing '%s'", path) : PF_OK;
}

int pf_write_stats_csv(const char* path, int append, int nx, int ny, double dt, int step_res_count, size_t n_rows, const pf_stats_row* rows)
{
    if(!path || (n_rows > 0 && !rows) || step_res_count < 0 || step_res_count > PF_MAX_STEP_RESIDUALS) return fail(nullptr, PF_ERR_INVALID, "pf_write_stats_csv: bad argument");
    FILE* f = fopen(path, append ? "ab" : "wb");
    if(!f) return fail(nullptr, PF_ERR_INVALID, "cannot open '%s': %s", path, strerror(errno));
    if(!append)
    {
        fprintf(f, "%i,%i,%lf\n", nx, ny, dt);
        fprintf(f, "\"time\",\"iter\",\"Phi_iters\",\"T_iters\",\"T_delta_L1\",\"T_delta_L2\",\"T_delta_max\",\"T_delta_min\","
                   "\"Phi_delta_L1\",\"Phi_delta_L2\",\"Phi_delta_max\",\"Phi_delta_min\"");
        for(int k = 0; k < step_res_count; k++)
            fprintf(f, ",\"step_res_L1[%i]\",\"step_res_L2[%i]\",\"step_res_max[%i]\",\"step_res_min[%i]\"", k, k, k, k);
        fprintf(f, "\n");
    }
    // every number goes through a 16-byte string, as float_vec_string_at / int_vec_string_at do (main.cpp:741-757)
    auto F = [&](double v) { char b[16] = {0}; snprintf(b, sizeof b, "%f", (float) v); fputs(b, f); };
    auto I = [&](long long v) { char b[16] = {0}; snprintf(b, sizeof b, "%lli", v); fputs(b, f); };
    for(size_t y = 0; y < n_rows; y++)
    {
        const pf_stats_row& r = rows[y];
        F(r.time); fputc(',', f); I(r.iter); fputc(',', f); I(r.Phi_iters); fputc(',', f); I(r.T_iters);
        const float v[8] = {r.T_delta_L1, r.T_delta_L2, r.T_delta_max, r.T_delta_min, r.Phi_delta_L1, r.Phi_delta_L2, r.Phi_delta_max, r.Phi_delta_min};
        for(int k = 0; k < 8; k++) { fputc(',', f); F(v[k]); }
        for(int k = 0; k < step_res_count; k++)
        {
            fputc(',', f); F(r.step_res_L1[k]); fputc(',', f); F(r.step_res_L2[k]);
            fputc(',', f); F(r.step_res_max[k]); fputc(',', f); F(r.step_res_min[k]);
        }
        fputc('\n', f);
    }
    int rc = ferror(f) ? -1 : 0;
    if(fclose(f) != 0) rc = -1;
    return rc ? fail(nullptr, PF_ERR_INVALID, "error writing '%s'", path) : PF_OK;
}

int pf_write_png(const char* path, int width, int height, const uint8_t* rgba, int flip_y)
{
    if(!path || !rgba || width <= 0 || height <= 0) return fail(nullptr, PF_ERR_INVALID, "pf_write_png: bad argument");
    FILE* f = fopen(path, "wb");
    if(!f) return fail(nullptr, PF_ERR_INVALID, "cannot open '%s': %s", path, strerror(errno));
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    fwrite(sig, 8, 1, f);
    uint8_t ihdr[13]; put_be32(ihdr, (uint32_t) width); put_be32(ihdr + 4, (uint32_t) height);
    ihdr[8] = 8; ihdr[9] = 6; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;           // 8-bit RGBA, no interlace
    int rc = png_chunk(f, "IHDR", ihdr, 13);
    // zlib stream of stored deflate blocks over the filtered scanlines (filter byte 0 + row)
    const size_t row = (size_t) width * 4 + 1, raw = row * (size_t) height;
    const size_t n_blocks = (raw + 65534) / 65535;
    std::string z; z.reserve(raw + n_blocks * 5 + 6);
    z.push_back((char) 0x78); z.push_back((char) 0x01);
    uint32_t a = 1, b = 0;                                                          // Adler-32 over the raw data
    std::string line; line.resize(row);
    size_t in_block = 0, left = raw;
    auto open_block = [&]() {
        const size_t n = left < 65535 ? left : 65535;
        z.push_back((char)(left <= 65535 ? 1 : 0));
        z.push_back((char)(n & 0xFF)); z.push_back((char)(n >> 8)); z.push_back((char)(~n & 0xFF)); z.push_back((char)((~n >> 8) & 0xFF));
        in_block = n;
    };
    for(int y = 0; y < height; y++)
    {
        const uint8_t* src = rgba + (size_t)(flip_y ? height - 1 - y : y) * (size_t) width * 4;
        line[0] = 0; memcpy(&line[1], src, (size_t) width * 4);
        for(size_t i = 0; i < row; i++)
        {
            if(in_block == 0) open_block();
            const uint8_t c = (uint8_t) line[i];
            z.push_back((char) c); in_block--; left--;
            a += c; if(a >= 65521) a -= 65521;
            b += a; if(b >= 65521) b -= 65521;
        }
    }
    uint8_t ad[4]; put_be32(ad, (b << 16) | a);
    z.append((const char*) ad, 4);
    if(rc == 0) rc = png_chunk(f, "IDAT", (const uint8_t*) z.data(), z.size());
    if(rc == 0) rc = png_chunk(f, "IEND", nullptr, 0);
    if(fclose(f) != 0) rc = -1;
    return rc ? fail(nullptr, PF_ERR_INVALID, "error writing '%s'", path) : PF_OK;
}

} // extern "C"
