// pic_state.inl -- host-side I/O around the engine, included at the end of
// pic_engine.cu:
//   * picb200_snapshot_png: the reference's image()/draw_particle()/save_png() (src/utility.cu:4-74) without
//     libpng -- a stored-deflate PNG encoder, CPU only, no graphics API;
//   * picb200_save_state / picb200_load_state: binary snapshot of the population and the run counters.  The
//     reference can only print the state as text (src/utility.cu:124-137, ~230 B per electron) and has no
//     restart; because the RNG is counter-based on (particle id, global step), a run resumed from a snapshot
//     continues bit for bit.

namespace picb200 {

static uint32_t crc32_update(uint32_t crc, const unsigned char* p, size_t n) {
    static uint32_t table[256];
    static bool ready = false;
    if (!ready) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            table[i] = c;
        }
        ready = true;
    }
    for (size_t i = 0; i < n; ++i) crc = table[(crc ^ p[i]) & 0xffu] ^ (crc >> 8);
    return crc;
}
static void put_be32(std::vector<unsigned char>& v, uint32_t x) {
    v.push_back((unsigned char)(x >> 24)); v.push_back((unsigned char)(x >> 16)); v.push_back((unsigned char)(x >> 8)); v.push_back((unsigned char)x);
}
static void png_chunk(std::vector<unsigned char>& out, const char* type, const std::vector<unsigned char>& data) {
    put_be32(out, (uint32_t)data.size());
    const size_t at = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), data.begin(), data.end());
    put_be32(out, crc32_update(0xffffffffu, out.data() + at, out.size() - at) ^ 0xffffffffu);
}
// 8-bit RGB image -> PNG with an uncompressed ("stored") zlib stream: valid for every PNG reader.
static bool write_png_rgb(const unsigned char* rgb, int width, int height, const char* path) {
    std::vector<unsigned char> raw;
    raw.reserve((size_t)height * ((size_t)width * 3 + 1));
    for (int y = 0; y < height; ++y) {
        raw.push_back(0);                                     // filter type 0 (None)
        raw.insert(raw.end(), rgb + (size_t)y * width * 3, rgb + (size_t)(y + 1) * width * 3);
    }
    std::vector<unsigned char> z;
    z.push_back(0x78); z.push_back(0x01);                    // zlib header: deflate, 32K window, no preset dictionary
    uint32_t a = 1, b = 0;                                    // Adler-32 of the raw stream
    for (unsigned char c : raw) { a = (a + c) % 65521u; b = (b + a) % 65521u; }
    for (size_t pos = 0; pos < raw.size();) {
        const size_t len = std::min<size_t>(65535, raw.size() - pos);
        z.push_back(pos + len == raw.size() ? 1 : 0);        // BFINAL, BTYPE = 00 (stored)
        z.push_back((unsigned char)(len & 0xff)); z.push_back((unsigned char)(len >> 8));
        z.push_back((unsigned char)(~len & 0xff)); z.push_back((unsigned char)((~len >> 8) & 0xff));
        z.insert(z.end(), raw.begin() + (long)pos, raw.begin() + (long)(pos + len));
        pos += len;
    }
    put_be32(z, (b << 16) | a);
    std::vector<unsigned char> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<unsigned char> ihdr;
    put_be32(ihdr, (uint32_t)width); put_be32(ihdr, (uint32_t)height);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);   // 8 bit, RGB, deflate, adaptive, no interlace
    png_chunk(out, "IHDR", ihdr);
    png_chunk(out, "IDAT", z);
    png_chunk(out, "IEND", {});
    FILE* fp = fopen(path, "wb");
    if (!fp) return false;
    const bool ok = fwrite(out.data(), 1, out.size(), fp) == out.size();
    fclose(fp);
    return ok;
}

struct StateHeader {
    char magic[8];                 // "PICB200S"
    uint32_t version, header_bytes;
    int64_t n;
    int32_t poisson_index, poisson_timestep;
    int64_t total_added, total_removed;
    uint64_t pushes, seed;
    int32_t sorted_by_id, reserved;
};
static_assert(sizeof(StateHeader) == 72, "state file header layout");

__global__ void k_restore_counters(DevState* st, int poisson_index, long long added, long long removed, unsigned long long pushes) {
    st->poisson_index = poisson_index;
    st->total_added = added;
    st->total_removed = removed;
    st->pushes = pushes;
}

}  // namespace picb200

// image() of src/utility.cu:52-74: 500 x 500 white canvas, an 11 x 11 black square per electron at pixel
// (x + i, 500 - (y + j)), i, j = -5..5, positions taken as pixel coordinates (a leftover of the reference's
// 2-D prototype: positions are metres in [0, 5.12], so every square lands in the bottom-left corner).
// The reference's bounds test lets row 500 and column 500 through (a write past the buffer); those pixels
// are skipped here.
extern "C" int picb200_snapshot_png(const float* x, const float* y, int64_t n, const char* path) {
    if (!path || (n > 0 && (!x || !y))) return fail("null argument", "picb200_snapshot_png");
    const int width = 500, height = 500;
    std::vector<unsigned char> img((size_t)width * height * 3, 255);
    for (int64_t e = 0; e < n; ++e) {
        const float px = x[e], py = y[e];
        for (int i = -5; i <= 5; ++i) {
            for (int j = -5; j <= 5; ++j) {
                if (py + j > height || py + j < 0 || px + i > width || px + i < 0) continue;      // src/utility.cu:41
                const int row = (int)(height - (py + j)), col = (int)(px + i);
                if (row < 0 || row >= height || col < 0 || col >= width) continue;
                unsigned char* p = &img[((size_t)row * width + col) * 3];
                p[0] = p[1] = p[2] = 0;
            }
        }
    }
    if (!write_png_rgb(img.data(), width, height, path)) return fail("cannot write the PNG file", "picb200_snapshot_png");
    return 0;
}

extern "C" int picb200_save_state(picb200_sim* sim, const char* path, int32_t sorted_by_id) {
    if (!sim || !path) return fail("null argument", "picb200_save_state");
    Sim& s = sim->s;
    if (s.sync_state()) return 1;
    const size_t n = (size_t)std::min(std::max(s.h_st->n, s.h_st->n_alloc), s.cfg.capacity);
    std::vector<float> f(n * 9 + 1);
    std::vector<int32_t> ts(n + 1);
    std::vector<uint64_t> id(n + 1);
    int got = 0;
    float* b = f.data();
    if (picb200_download(sim, &got, b, b + n, b + 2 * n, b + 3 * n, b + 4 * n, b + 5 * n, b + 6 * n, b + 7 * n, b + 8 * n, ts.data(), id.data())) return 1;
    std::vector<uint32_t> perm(n);
    for (size_t i = 0; i < n; ++i) perm[i] = (uint32_t)i;
    if (sorted_by_id) std::sort(perm.begin(), perm.end(), [&](uint32_t p, uint32_t q) { return id[p] < id[q]; });   // slot order is a race; id order is not
    StateHeader h{};
    memcpy(h.magic, "PICB200S", 8);
    h.version = 1; h.header_bytes = (uint32_t)sizeof(StateHeader);
    h.n = (int64_t)n;
    h.poisson_index = s.h_st->poisson_index; h.poisson_timestep = s.cfg.poisson_timestep;
    h.total_added = s.h_st->total_added; h.total_removed = s.h_st->total_removed;
    h.pushes = s.h_st->pushes; h.seed = s.cfg.seed;
    h.sorted_by_id = sorted_by_id ? 1 : 0;
    FILE* fp = fopen(path, "wb");
    if (!fp) return fail("cannot open the state file for writing", "picb200_save_state");
    bool ok = fwrite(&h, sizeof h, 1, fp) == 1;
    std::vector<float> col(n + 1);
    for (int k = 0; k < 9 && ok; ++k) {
        for (size_t i = 0; i < n; ++i) col[i] = b[(size_t)k * n + perm[i]];
        ok = fwrite(col.data(), 4, n, fp) == n;
    }
    if (ok) {
        std::vector<int32_t> t2(n + 1);
        for (size_t i = 0; i < n; ++i) t2[i] = ts[perm[i]];
        ok = fwrite(t2.data(), 4, n, fp) == n;
    }
    if (ok) {
        std::vector<uint64_t> i2(n + 1);
        for (size_t i = 0; i < n; ++i) i2[i] = id[perm[i]];
        ok = fwrite(i2.data(), 8, n, fp) == n;
    }
    fclose(fp);
    return ok ? 0 : fail("short write", "picb200_save_state");
}

extern "C" int picb200_load_state(picb200_sim* sim, const char* path) {
    if (!sim || !path) return fail("null argument", "picb200_load_state");
    Sim& s = sim->s;
    FILE* fp = fopen(path, "rb");
    if (!fp) return fail("cannot open the state file", "picb200_load_state");
    StateHeader h{};
    bool ok = fread(&h, sizeof h, 1, fp) == 1 && !memcmp(h.magic, "PICB200S", 8) && h.version == 1 && h.header_bytes == sizeof(StateHeader);
    if (!ok) { fclose(fp); return fail("not a picb200 state file", "picb200_load_state"); }
    if (h.n < 0 || h.n > (int64_t)s.cfg.capacity) { fclose(fp); return fail("the snapshot holds more electrons than this sim's capacity", "picb200_load_state"); }
    if (h.poisson_timestep != s.cfg.poisson_timestep || h.seed != s.cfg.seed) { fclose(fp); return fail("the snapshot was taken with another poisson_timestep or seed", "picb200_load_state"); }
    const size_t n = (size_t)h.n;
    std::vector<float> f(n * 9 + 1);
    std::vector<int32_t> ts(n + 1);
    std::vector<uint64_t> id(n + 1);
    ok = fread(f.data(), 4, n * 9, fp) == n * 9 && fread(ts.data(), 4, n, fp) == n && fread(id.data(), 8, n, fp) == n;
    fclose(fp);
    if (!ok) return fail("truncated state file", "picb200_load_state");
    float* b = f.data();
    if (picb200_upload(sim, (int32_t)n, b, b + n, b + 2 * n, b + 3 * n, b + 4 * n, b + 5 * n, b + 6 * n, b + 7 * n, b + 8 * n, ts.data(), id.data())) return 1;
    k_restore_counters<<<1, 1, 0, s.stream>>>(s.st, h.poisson_index, (long long)h.total_added, (long long)h.total_removed, (unsigned long long)h.pushes);
    CK_LAST("State restore");
    return s.sync_state();
}
