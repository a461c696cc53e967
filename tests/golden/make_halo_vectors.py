#!/usr/bin/env python
"""Golden vectors for the halo / tripole-fold code from the REFERENCE'S OWN TEXT (build container only).

Executed mechanically by tests/golden/fortran_exec.py, straight from /root/reference, MPI build
(the one COSIMA runs; on one task every copy is a "local copy" and the send / receive loops are empty):

  source/ice_blocks.F90    create_blocks, ice_blocksGetNbrID, get_block_parameter
  mpi/ice_boundary.F90     ice_HaloCreate, ice_HaloMsgCreate, ice_HaloIncrementMsgCount,
                           ice_HaloUpdate2DR8, ice_HaloUpdate3DR8, ice_HaloUpdate2DI4, ice_HaloUpdate_stress

Only ice_distributionGet / ice_distributionGetBlockLoc (one task owning every block, local id = block
id) and global_maxval (identity on one task) are stood in for.  For each domain (size, block size,
boundary types) the halo structure is created by the reference's code and every update the EVP driver
issues (source/ice_dyn_evp.F90:210, 311-315, 394-398, 425-480; source/ice_grid.F90 t2ugrid_vector /
u2tgrid_vector) is run on seeded random blocks whose ghost cells hold garbage on entry.  Inputs, outputs
and the blocks' global index maps go to tests/golden/ref_halo_vectors.npz;
tests/test_oracle_reference_text.py demands the same cells from the C oracle, bit for bit.

    python tests/golden/make_halo_vectors.py
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import fortran_exec as fe  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(HERE, "ref_halo_vectors.npz")

# name: (nx_global, ny_global, block_size_x, block_size_y, ew_boundary_type, ns_boundary_type)
DOMAINS = {
    "tripole_1x1": (16, 10, 16, 10, "cyclic", "tripole"),
    "tripole_2x2": (16, 10, 8, 5, "cyclic", "tripole"),
    "tripole_4x1": (24, 7, 6, 7, "cyclic", "tripole"),       # x-slabs, the product's own decomposition
    "open_1x1": (12, 8, 12, 8, "cyclic", "open"),            # gx3 / gx1: cyclic east-west, open north-south
    "open_3x2": (12, 8, 4, 4, "cyclic", "open"),
    "closed_1x1": (10, 6, 10, 6, "closed", "closed"),
    "closed_2x1": (10, 6, 5, 6, "closed", "closed"),
}
LOCS = {"center": "field_loc_center", "necorner": "field_loc_necorner"}
KINDS = {"scalar": "field_type_scalar", "vector": "field_type_vector"}
GARBAGE = 7.0e33


def _lines(rel, defined=()):
    return fe.preprocess(open(os.path.join(REF, rel)).read(), defined)


def machine(nxg, nyg, bsx, bsy, ew, ns):
    """A fresh copy of the reference's block + halo modules for one domain; returns (module, halo)."""
    nsd = {"block_size_x": bsx, "block_size_y": bsy}
    fe.parse_parameters(_lines("drivers/cice/ice_constants.F90"), nsd)
    blocks, bnd = _lines("source/ice_blocks.F90"), _lines("mpi/ice_boundary.F90")
    fe.parse_parameters(blocks, nsd)                 # nghost, nx_block, ny_block, ice_blocksNorth ...
    fe.parse_parameters(bnd, nsd)
    m = fe.Module(nsd)

    @fe.fstub(["dist", "nprocs", "communicator", "numlocalblocks"], out=["nprocs", "communicator", "numlocalblocks"])
    def ice_distributionget(dist, nprocs=None, communicator=None, numlocalblocks=None):
        return {"nprocs": 1, "communicator": 0, "numlocalblocks": m.ns["nblocks_tot"]}

    @fe.fstub(["dist", "blockid", "processor", "localid"], out=["processor", "localid"])
    def ice_distributiongetblockloc(dist, blockid, processor, localid):
        return {"processor": 1, "localid": blockid}

    m.ns.update(ice_distributionget=ice_distributionget, ice_distributiongetblockloc=ice_distributiongetblockloc,
                global_maxval=lambda x, dist: x, my_task=0, master_task=0, nu_diag=6, mpi_status_size=6,
                # module variables of ice_blocks / ice_boundary, unallocated
                all_blocks=None, all_blocks_ij=None, i_global=None, j_global=None, nblocks_x=0, nblocks_y=0,
                nblocks_tot=0, buftripoler8=None, buftripolei4=None, buftripoler4=None, bufsizesend=0, bufsizerecv=0,
                bufsendr8=None, bufrecvr8=None, bufsendi4=None, bufrecvi4=None, bufsendr4=None, bufrecvr4=None)
    m.load(blocks, ["get_block_parameter", "create_blocks", "ice_blocksgetnbrid"], derived_arrays=["all_blocks"])
    m.load(bnd, ["ice_haloincrementmsgcount", "ice_halomsgcreate", "ice_halocreate", "ice_haloupdate2dr8",
                 "ice_haloupdate3dr8", "ice_haloupdate2di4", "ice_haloupdate_stress"])
    m.call("create_blocks", nx_global=nxg, ny_global=nyg, ew_boundary_type=ew, ns_boundary_type=ns)
    halo = m.call("ice_halocreate", dist=None, nsboundarytype=ns, ewboundarytype=ew, nxglobal=nxg)
    return m, halo


def index_maps(m):
    """(nblocks, nx_block) i_glob and (nblocks, ny_block) j_glob as create_blocks left them, + (jblock, nblocks_y)."""
    nb = m.ns["nblocks_tot"]
    ig = np.array([[m.ns["all_blocks"](n).i_glob(i) for i in range(1, m.ns["nx_block"] + 1)] for n in range(1, nb + 1)])
    jg = np.array([[m.ns["all_blocks"](n).j_glob(j) for j in range(1, m.ns["ny_block"] + 1)] for n in range(1, nb + 1)])
    jb = np.array([m.ns["all_blocks"](n).jblock for n in range(1, nb + 1)])
    return ig.astype(np.int64), jg.astype(np.int64), jb.astype(np.int64)


def scatter(m, glob, garbage, dtype=float):
    """Global physical field (ny, nx) -> Fortran (nx_block, ny_block, nblocks); ghosts = garbage."""
    nxb, nyb, nb = m.ns["nx_block"], m.ns["ny_block"], m.ns["nblocks_tot"]
    arr = fe.FArr([(1, nxb), (1, nyb), (1, nb)], dtype=dtype, fill=garbage)
    for n in range(1, nb + 1):
        b = m.ns["all_blocks"](n)
        for j in range(b.jlo, b.jhi + 1):
            for i in range(b.ilo, b.ihi + 1):
                arr[i, j, n] = glob[b.j_glob(j) - 1, b.i_glob(i) - 1]
    return arr


def blocks_of(arr):
    """Fortran (nx_block, ny_block, nblocks) -> numpy (nblocks, ny_block, nx_block)."""
    return np.ascontiguousarray(np.transpose(arr.a, (2, 1, 0)))


def main(out_path=OUT):
    pack = {}
    for dname, (nxg, nyg, bsx, bsy, ew, ns) in DOMAINS.items():
        m, halo = machine(nxg, nyg, bsx, bsy, ew, ns)
        ig, jg, jb = index_maps(m)
        pack[f"{dname}__dims"] = np.array([nxg, nyg, bsx, bsy, m.ns["nblocks_x"], m.ns["nblocks_y"]], np.int64)
        pack[f"{dname}__bnd"] = np.array([ew, ns])
        pack[f"{dname}__i_glob"], pack[f"{dname}__j_glob"], pack[f"{dname}__jblock"] = ig, jg, jb
        pack[f"{dname}__copies"] = np.int64(halo.numlocalcopies)
        rng = np.random.default_rng(sum(map(ord, dname)))
        # --- 2-D double precision, every loc x kind the path uses, with and without fillValue
        for lname, loc in LOCS.items():
            for kname, kind in KINDS.items():
                for fill in (None, -3.5):
                    glob = rng.uniform(-1.0, 1.0, (nyg, nxg))
                    arr = scatter(m, glob, GARBAGE)
                    kw = {} if fill is None else {"fillvalue": fill}
                    m.call("ice_haloupdate2dr8", array=arr, halo=halo, fieldloc=m.ns[loc], fieldkind=m.ns[kind], **kw)
                    key = f"{dname}__r8__{lname}__{kname}__{'nofill' if fill is None else 'fill'}"
                    pack[key + "__in"], pack[key + "__out"] = glob, blocks_of(arr)
        # --- the velocity update of the subcycle loop: fld2(:,:,1:2,:) through the 3-D routine
        u, v = rng.uniform(-1.0, 1.0, (nyg, nxg)), rng.uniform(-1.0, 1.0, (nyg, nxg))
        a_u, a_v = scatter(m, u, GARBAGE), scatter(m, v, GARBAGE)
        fld2 = fe.FArr([(1, m.ns["nx_block"]), (1, m.ns["ny_block"]), (1, 2), (1, m.ns["nblocks_tot"])])
        fld2.a[:, :, 0, :], fld2.a[:, :, 1, :] = a_u.a, a_v.a
        m.call("ice_haloupdate3dr8", array=fld2, halo=halo, fieldloc=m.ns["field_loc_necorner"],
               fieldkind=m.ns["field_type_vector"])
        pack[f"{dname}__fld2__in"] = np.stack([u, v])
        pack[f"{dname}__fld2__out"] = np.ascontiguousarray(np.transpose(fld2.a, (2, 3, 1, 0)))   # (2, nblocks, ny_block, nx_block)
        # --- icetmask: 2-D integer, center scalar
        msk = (rng.uniform(size=(nyg, nxg)) > 0.4).astype(np.int64)
        arr = scatter(m, msk, 77, dtype=int)
        m.call("ice_haloupdate2di4", array=arr, halo=halo, fieldloc=m.ns["field_loc_center"],
               fieldkind=m.ns["field_type_scalar"])
        pack[f"{dname}__i4__in"], pack[f"{dname}__i4__out"] = msk.astype(np.int32), blocks_of(arr).astype(np.int32)
        # --- ice_HaloUpdate_stress(array1, array2): ghosts keep their values where nothing is copied
        g1, g2 = rng.uniform(-1.0, 1.0, (nyg, nxg)), rng.uniform(-1.0, 1.0, (nyg, nxg))
        a1, a2 = scatter(m, g1, GARBAGE), scatter(m, g2, GARBAGE)
        ghosts_in = rng.uniform(2.0, 3.0, a1.a.shape)              # recognisable ghost values on entry
        phys = a1.a != GARBAGE
        a1.a[~phys] = ghosts_in[~phys]
        pack[f"{dname}__stress__a1_blocks_in"] = blocks_of(a1)
        m.call("ice_haloupdate_stress", array1=a1, array2=a2, halo=halo, fieldloc=m.ns["field_loc_center"],
               fieldkind=m.ns["field_type_scalar"])
        pack[f"{dname}__stress__in1"], pack[f"{dname}__stress__in2"] = g1, g2
        pack[f"{dname}__stress__out"] = blocks_of(a1)
    if fe.UNTRANSLATED:
        print("statements outside the interpreter's subset (raise if reached):")
        for u_ in sorted(set(fe.UNTRANSLATED)):
            print("   ", u_)
    np.savez_compressed(out_path, **pack)
    print(f"{out_path}: {len(pack)} arrays, {os.path.getsize(out_path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
