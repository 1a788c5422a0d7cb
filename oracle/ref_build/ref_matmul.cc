// TEST INFRASTRUCTURE (oracle/): drives the reference's own Math::Matrix<f32> * Math::Vector<f32> -- the product
// signal-matrix-multiplication-f32 evaluates (src/Signal/MatrixMult.hh, src/Math/Matrix.hh:482-489) -- compiled with
// the reference's flags, to pin oracle/gmm_oracle.c orc_lda and the golden fixture tests/golden/lda_matmul.npz.
// usage: ref_matmul <in> <out>;  in = u32 rows, cols, n; f32 matrix[rows*cols]; f32 vectors[n*cols];  out = f32 [n*rows]
#include <Math/Matrix.hh>
#include <Math/Vector.hh>
#include <cstdio>
#include <vector>

int main(int argc, char** argv) {
    if (argc != 3)
        return 2;
    FILE* f = fopen(argv[1], "rb");
    if (!f)
        return 3;
    u32 hdr[3];
    if (fread(hdr, 4, 3, f) != 3)
        return 4;
    const u32         rows = hdr[0], cols = hdr[1], n = hdr[2];
    std::vector<f32>  m((size_t)rows * cols), v((size_t)n * cols), out((size_t)n * rows);
    if (fread(m.data(), 4, m.size(), f) != m.size() || fread(v.data(), 4, v.size(), f) != v.size())
        return 5;
    fclose(f);
    Math::Matrix<f32> matrix(rows, cols);
    for (u32 r = 0; r < rows; ++r)
        for (u32 c = 0; c < cols; ++c)
            matrix[r][c] = m[(size_t)r * cols + c];
    for (u32 i = 0; i < n; ++i) {
        Math::Vector<f32> in(cols);
        for (u32 c = 0; c < cols; ++c)
            in[c] = v[(size_t)i * cols + c];
        Math::Vector<f32> res = matrix * in;
        for (u32 r = 0; r < rows; ++r)
            out[(size_t)i * rows + r] = res[r];
    }
    f = fopen(argv[2], "wb");
    if (!f || fwrite(out.data(), 4, out.size(), f) != out.size())
        return 6;
    fclose(f);
    return 0;
}
