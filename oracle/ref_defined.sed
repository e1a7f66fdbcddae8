# oracle/ref_defined.sed -- TEST INFRASTRUCTURE. The REF_DEFINED patch list of SURVEY.md 8(c), applied to the
# reference's rpf.cpp as it streams through sed at build time (addresses are reference line numbers).
# P1 (rpf.cpp:368,440,464): accumulate D_c_fk[j] = sum_i MI(c_i,f_j) and use it for W_c_fk instead of the
#     out-of-bounds D_f_ck[i].
368s|^| SampleF D_c_fk; for (auto \&v_ : D_c_fk) v_ = 0;|
440s|D_f_ck\[i\] += MutualInformation(colors_data\[i\], features_data\[j\]);|{ double m_ = MutualInformation(colors_data[i], features_data[j]); D_f_ck[i] += m_; D_c_fk[j] += m_; }|
# P2 (rpf.cpp:464,465,470): 0/0 -> 0, the idiom of ops.h:48
464s|W_c_fk\[i\] = D_f_ck\[i\] / (D_f_c + D_r_c + D_p_c);|W_c_fk[i] = REF_RATIO(D_c_fk[i], (D_f_c + D_r_c + D_p_c));|
465s|W_r_fk\[i\] = D_r_fk\[i\] / (D_r_fk\[i\] + D_p_fk\[i\]);|W_r_fk[i] = REF_RATIO(D_r_fk[i], (D_r_fk[i] + D_p_fk[i]));|
470s|W_r_ck\[i\] = D_r_ck\[i\] / (D_r_ck\[i\] + D_p_ck\[i\]);|W_r_ck[i] = REF_RATIO(D_r_ck[i], (D_r_ck[i] + D_p_ck[i]));|
# lifted literal (rpf.cpp:533)
533s|double sigma_fc_seed = 0.002;|double sigma_fc_seed = ::pbrt::ref_sigma_fc_seed;|
# P3 (rpf.cpp:601): a variance that rounded below zero gives sigma 0, not NaN (neighbourhood statistics only)
601s|$| for (auto \&v_ : stdDev) if (v_ != v_) v_ = 0;|
302,733p
