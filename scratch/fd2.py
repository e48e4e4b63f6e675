import sys
exec(open('/root/repo/scratch/fd_adjoint.py').read().split("run('llvm_ad_mono'")[0])
def run2(variant, integ, phase, alb, st, spp=65536):
    f = homog(variant, integ, spp=spp, phase=phase, albedo=alb, sigma_t=st)
    o = orc.OracleScene(f)
    o.render_1_backward(7, spp, [1.0,0,0,0], nthreads=16)
    gs = o.grad_get_scalar(0)
    for key, idx, h in (('albedo', 0, 1e-2), ('sigma_t', 3, 3e-2)):
        vals = []
        for sgn in (+1, -1):
            f2 = homog(variant, integ, spp=spp, phase=phase, albedo=alb + (sgn*h if key=='albedo' else 0), sigma_t=st + (sgn*h if key=='sigma_t' else 0))
            vals.append(orc.OracleScene(f2).render_1(11 if key=='sigma_t' else 7, spp, nthreads=16)[0])
        fd = (vals[0]-vals[1])/(2*h)
        print(variant, integ, key, 'FD', fd, 'adj', gs[idx], 'ratio', gs[idx]/fd)
run2('llvm_ad_mono_polarized', 'prb_volpathaos', {'type':'rayleigh'}, 0.8, 0.4)
