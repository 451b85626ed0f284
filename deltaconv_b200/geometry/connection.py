"""Parallel-transport helpers (reference: deltaconv/geometry/connection.py).  Not on the hot path
(no model calls them; `rotate_around` is used by tests) — plain device-agnostic torch
expressions kept for API parity."""
import torch

from .utils import batch_dot

__all__ = ['build_transport', 'angle_in_plane', 'rotate_around']


def _unit(a, floor):
    return a / torch.linalg.norm(a, dim=-1, keepdim=True).clamp(floor)


def angle_in_plane(u, v, normal):
    u_flat = _unit(u - batch_dot(u, normal) * normal, 1e-8)
    side = _unit(torch.linalg.cross(normal, u_flat, dim=-1), 1e-8)
    return torch.atan2(batch_dot(v, side), batch_dot(v, u_flat))


def rotate_around(v, axis, angle):
    if angle.dim() == 1:
        angle = angle.unsqueeze(-1)
    along = axis * batch_dot(v, axis)
    across = v - along
    length = torch.linalg.norm(across, dim=-1, keepdim=True).clamp(1e-8)
    e0 = across / length
    e1 = torch.linalg.cross(axis, e0, dim=-1)
    turned = length * (torch.cos(angle) * e0 + torch.sin(angle) * e1) + along
    return torch.where(length > 0, turned, along)


def build_transport(target_n, target_x, target_y, source_n, source_x, non_oriented=True):
    flipped = batch_dot(source_n, target_n) < 0
    target_n = torch.where(flipped, -target_n, target_n)
    target_y = torch.where(flipped, -target_y, target_y)
    axis = torch.linalg.cross(target_n, source_n, dim=-1)
    axis_len = torch.linalg.norm(axis, dim=-1, keepdim=True)
    axis = torch.where(axis_len > 1e-6, axis / axis_len, source_x)
    moved = rotate_around(source_x, axis, angle_in_plane(source_n, target_n, axis))
    comp = torch.cat([batch_dot(moved, target_x), batch_dot(moved, target_y)], dim=1)
    comp_len = torch.linalg.norm(comp, dim=-1, keepdim=True)
    unit_x = torch.zeros_like(comp)
    unit_x[:, 0] = 1
    comp = torch.where(comp_len > 1e-6, comp / comp_len, unit_x)
    sign = torch.ones(comp.size(0), device=comp.device)
    if non_oriented:
        sign = torch.where(flipped.flatten(), -sign, sign)
    return torch.stack([comp[:, 0], -comp[:, 1], comp[:, 1] * sign, comp[:, 0] * sign], dim=1)
