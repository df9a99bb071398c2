"""paddle.linalg (reference: python/paddle/tensor/linalg.py) on torch.linalg; Paddle argument names kept."""
from __future__ import annotations

import torch


def norm(x, p=None, axis=None, keepdim=False, name=None):
    if p is None or p == "fro":
        return torch.linalg.norm(x, dim=axis, keepdim=keepdim) if axis is not None else torch.linalg.norm(x)
    if p == "nuc":
        return torch.linalg.matrix_norm(x, ord="nuc", dim=tuple(axis) if axis is not None else (-2, -1), keepdim=keepdim)
    return torch.linalg.vector_norm(x, ord=p, dim=axis, keepdim=keepdim)


def vector_norm(x, p=2.0, axis=None, keepdim=False, name=None):
    return torch.linalg.vector_norm(x, ord=p, dim=axis, keepdim=keepdim)


def matrix_norm(x, p="fro", axis=(-2, -1), keepdim=False, name=None):
    return torch.linalg.matrix_norm(x, ord=p, dim=tuple(axis), keepdim=keepdim)


def inv(x, name=None):
    return torch.linalg.inv(x)


def pinv(x, rcond=1e-15, hermitian=False, name=None):
    return torch.linalg.pinv(x, rtol=rcond, hermitian=hermitian)


def det(x, name=None):
    return torch.linalg.det(x)


def slogdet(x, name=None):
    s, l = torch.linalg.slogdet(x)
    return torch.stack([s, l])


def svd(x, full_matrices=False, name=None):
    u, s, vh = torch.linalg.svd(x, full_matrices=full_matrices)
    return u, s, vh


def qr(x, mode="reduced", name=None):
    return torch.linalg.qr(x, mode=mode)


def cholesky(x, upper=False, name=None):
    return torch.linalg.cholesky(x, upper=upper)


def cholesky_solve(x, y, upper=False, name=None):
    return torch.cholesky_solve(x, y, upper=upper)


def eig(x, name=None):
    return torch.linalg.eig(x)


def eigh(x, UPLO="L", name=None):
    return torch.linalg.eigh(x, UPLO=UPLO)


def eigvals(x, name=None):
    return torch.linalg.eigvals(x)


def eigvalsh(x, UPLO="L", name=None):
    return torch.linalg.eigvalsh(x, UPLO=UPLO)


def solve(x, y, name=None):
    return torch.linalg.solve(x, y)


def triangular_solve(x, y, upper=True, transpose=False, unitriangular=False, name=None):
    a = x.transpose(-1, -2) if transpose else x
    return torch.linalg.solve_triangular(a, y, upper=upper != transpose, unitriangular=unitriangular)


def lstsq(x, y, rcond=None, driver=None, name=None):
    r = torch.linalg.lstsq(x, y, rcond=rcond, driver=driver)
    return r.solution, r.residuals, r.rank, r.singular_values


def lu(x, pivot=True, get_infos=False, name=None):
    lu_, piv, info = torch.linalg.lu_factor_ex(x, pivot=pivot)
    return (lu_, piv + 0, info) if get_infos else (lu_, piv + 0)


def matrix_power(x, n, name=None):
    return torch.linalg.matrix_power(x, n)


def matrix_rank(x, tol=None, hermitian=False, name=None):
    return torch.linalg.matrix_rank(x, tol=tol, hermitian=hermitian)


def cond(x, p=None, name=None):
    return torch.linalg.cond(x, p=p)


def multi_dot(x, name=None):
    return torch.linalg.multi_dot(list(x))


def cov(x, rowvar=True, ddof=True, fweights=None, aweights=None, name=None):
    return torch.cov(x if rowvar else x.t(), correction=int(ddof), fweights=fweights, aweights=aweights)


def corrcoef(x, rowvar=True, name=None):
    return torch.corrcoef(x if rowvar else x.t())


def matrix_exp(x, name=None):
    return torch.linalg.matrix_exp(x)


def householder_product(x, tau, name=None):
    return torch.linalg.householder_product(x, tau)


def cholesky_inverse(x, upper=False, name=None):
    return torch.cholesky_inverse(x, upper=upper)


def svd_lowrank(x, q=None, niter=2, M=None, name=None):
    return torch.svd_lowrank(x, q=q if q is not None else min(6, *x.shape[-2:]), niter=niter, M=M)


def pca_lowrank(x, q=None, center=True, niter=2, name=None):
    return torch.pca_lowrank(x, q=q, center=center, niter=niter)


def lu_unpack(x, y, unpack_ludata=True, unpack_pivots=True, name=None):
    return torch.lu_unpack(x, y.to(torch.int32), unpack_data=unpack_ludata, unpack_pivots=unpack_pivots)


def ormqr(x, tau, y, left=True, transpose=False, name=None):
    return torch.ormqr(x, tau, y, left=left, transpose=transpose)
