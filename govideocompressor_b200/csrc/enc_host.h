// enc_host.h -- host-side helpers of the encoder kernels (pure C++): quantiser constants, geometry.
#pragma once
#include "enc_types.h"

namespace gvc {

// x264 encoder/set.c semantics for flat matrices (see enc_types.h QuantParams); intra selects the dead zone.
void fill_quant_params(QuantParams &q, int qp, bool intra);
void fill_quant8_params(Quant8Params &q, int qp, bool intra);
int  chroma_qp(int qp);     // H.264 Table 8-15, chroma_qp_index_offset = 0
int  lambda_for_qp(int qp); // x264_lambda_tab

inline int align_up(int v, int a) { return (v + a - 1) / a * a; }
EncGeom make_geom(int w, int h);

}  // namespace gvc
