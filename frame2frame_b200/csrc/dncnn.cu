// dncnn.cu -- placeholder translation unit (kernels land here next)
namespace f2f { long g_dncnn_launches = 0; }
