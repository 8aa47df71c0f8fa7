// TEST INFRASTRUCTURE ONLY — supplied stand-in for the reference's EMPTY Kernel/X2inv.cl.
//
// The reference kernel indexes X2inv[(uint) n] inside superSamplingStopCriteria
// (Kernel/PathTracer_FullKernel.cl:1168), which is outside every #if, so the translation unit
// needs a declaration to compile at all. The function is only CALLED under
// `#if SUPER_SAMPLING`, which no configuration in this repo enables, so the values below are
// never read. They are the 95 % quantile of a chi-square law with n degrees of freedom via the
// Wilson–Hilferty approximation, n = 0..63; anything else would do equally well.
static const float X2inv[64] = {
    1.0f, 3.8415f, 5.9915f, 7.8147f, 9.4877f, 11.0705f, 12.5916f, 14.0671f,
    15.5073f, 16.9190f, 18.3070f, 19.6751f, 21.0261f, 22.3620f, 23.6848f, 24.9958f,
    26.2962f, 27.5871f, 28.8693f, 30.1435f, 31.4104f, 32.6706f, 33.9244f, 35.1725f,
    36.4150f, 37.6525f, 38.8851f, 40.1133f, 41.3371f, 42.5570f, 43.7730f, 44.9853f,
    46.1943f, 47.3999f, 48.6024f, 49.8018f, 50.9985f, 52.1923f, 53.3835f, 54.5722f,
    55.7585f, 56.9424f, 58.1240f, 59.3035f, 60.4809f, 61.6562f, 62.8296f, 64.0011f,
    65.1708f, 66.3386f, 67.5048f, 68.6693f, 69.8322f, 70.9935f, 72.1532f, 73.3115f,
    74.4683f, 75.6237f, 76.7778f, 77.9305f, 79.0819f, 80.2321f, 81.3810f, 82.5287f};
