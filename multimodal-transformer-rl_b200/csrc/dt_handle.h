// The Decision-Transformer handle behind `dtb_dt` (C ABI) and the order of the entries in its flat parameter buffer; shared by
// model.cu (training / full forward) and decode.cu (KV-cached get_action).
#pragma once
#include "../../include/dtb200.h"
#include "kernels.h"
#include "side.h"

#include <vector>

namespace dtb {
enum { E_TABLE = 0, E_WR, E_BR, E_WS, E_BS, E_WA, E_BA, E_LNW, E_LNB, E_COUNT };
enum { L_LN1W = 0, L_LN1B, L_WQKV, L_BQKV, L_WO, L_BO, L_LN2W, L_LN2B, L_W1, L_B1, L_W2, L_B2, L_COUNT };
enum { T_LNFW = 0, T_LNFB, T_WPS, T_WPR, T_WPA, T_BPS, T_BPR, T_BPA, T_COUNT };

}  // namespace dtb

struct dtb_dt {
  dtb_dt_config c;
  int n_entries;
  std::vector<long long> off, num;
  long long total;
  int sd_small;  // fused fp32 state projection (sd <= 64) vs tensor-core projection of a wide state
  int P8, ZA8, ZS8;
  mutable dtb::SideStream side;  // weight-gradient branch of the backward
  // dropout (training mode): probabilities of the three GPT-2 sites and the device-resident generator state {seed_lo, seed_hi, step}
  float p_embd = 0.f, p_attn = 0.f, p_resid = 0.f;
  const void* drop_state = nullptr;
  dtb::DropSite drop_embd() const { return dtb::drop_make(drop_state, 0u, p_embd); }
  dtb::DropSite drop_attn(int l) const { return dtb::drop_make(drop_state, 1u + 3u * (uint32_t)l, p_attn); }
  dtb::DropSite drop_resid(int l, int which) const { return dtb::drop_make(drop_state, 2u + (uint32_t)which + 3u * (uint32_t)l, p_resid); }
  long long e(int id) const { return off[id]; }
  long long l(int layer, int id) const { return off[dtb::E_COUNT + layer * dtb::L_COUNT + id]; }
  long long t(int id) const { return off[dtb::E_COUNT + c.n_layer * dtb::L_COUNT + id]; }
};

